"""PSBROWS1 rows files (psb200/data/rows_file.py, tools/csv_to_rows.py): a CSV converted once into the palette wire
format streams through the same producer thread (CsvProducer.java:36-87) with nothing to tokenise, and gives the same
rows, labels, partitions and packed batches as the CSV it came from."""
import os
import subprocess
import sys
import time

import numpy as np
import pytest

from psb200.data import ReviewGenerator, is_rows_file, read_csv, read_rows_file, write_csv, write_rows_file
from psb200.ops import native
from psb200.utils.logs import parse_log

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not native.available(), reason="libpsb200.so not built")


def _drain(prod, packed=False, timeout=30.0):
    rows, labels, workers, eof, t0 = [], [], [], 0, time.time()
    while time.time() - t0 < timeout:
        b, n, eof = prod.poll()
        if b is None:
            if eof:
                break
            time.sleep(0.001)
            continue
        rows.append(prod.packed(b, n).to_dense() if packed else prod.rows[b, :n].numpy().copy())
        labels.append(prod.labels[b, :n].numpy().copy()), workers.append(prod.workers[b, :n].numpy().copy())
        prod.release(b)
    cat = lambda xs: np.concatenate(xs) if xs else np.zeros(0)
    return cat(rows), cat(labels), cat(workers), eof


@pytest.fixture(scope="module")
def files(tmp_path_factory):
    d = tmp_path_factory.mktemp("rows")
    gen = ReviewGenerator(num_features=96, vocab=900, mean_len=25)
    batch = gen.train_split(1500)
    batch.values[: batch.indptr[3]] = np.random.default_rng(0).standard_normal(batch.indptr[3])   # > 16 distinct values: raw rows
    csv, rows = str(d / "t.csv"), str(d / "t.psbrows")
    write_csv(csv, batch)
    text = read_csv(csv)                                          # the values as written to the CSV
    size = write_rows_file(rows, text, chunk=400)
    assert size == os.path.getsize(rows)
    return csv, rows, text


def test_round_trip_and_size(files):
    csv, rows, text = files
    assert is_rows_file(rows) and not is_rows_file(csv) and not is_rows_file(rows + ".missing")
    back = read_rows_file(rows)
    assert len(back) == len(text) and back.num_features == 96 and np.array_equal(back.labels, text.labels)
    assert np.array_equal(back.to_dense(np.float16).view(np.uint16), text.to_dense(np.float16).view(np.uint16))
    assert os.path.getsize(rows) < 0.25 * os.path.getsize(csv)
    with pytest.raises(ValueError):
        read_rows_file(csv)


@pytest.mark.parametrize("packed", [False, True, "palette"])
@pytest.mark.parametrize("threads", ["1", "3"])
def test_producer_streams_a_rows_file_like_the_csv_it_came_from(files, monkeypatch, packed, threads):
    csv, rows, text = files
    monkeypatch.setenv("PSB_PRODUCER_THREADS", threads)
    out = {}
    for path in (csv, rows):
        prod = native.StreamProducer(path, 4, 1, 2, 96, 0, batch_rows=128, num_batches=3, packed=packed)   # partitions 1, 2
        out[path] = _drain(prod, packed=bool(packed))
        prod.stop()
    (r0, l0, w0, e0), (r1, l1, w1, e1) = out[csv], out[rows]
    idx = np.nonzero(np.isin(np.arange(len(text)) % 4, (1, 2)))[0]
    assert e0 == e1 == 1 and len(r1) == len(idx)
    assert np.array_equal(l1, l0) and np.array_equal(w1, w0) and np.array_equal(w1, idx % 4 - 1)
    assert np.array_equal(r1.view(np.uint16), r0.view(np.uint16))
    assert np.array_equal(r1.view(np.uint16), text.to_dense(np.float16)[idx].view(np.uint16))


def test_damaged_or_mismatched_rows_files_are_refused(files, tmp_path):
    csv, rows, text = files
    raw = open(rows, "rb").read()
    cut = str(tmp_path / "cut.psbrows")
    open(cut, "wb").write(raw[: len(raw) - 100])                   # payload shorter than the header says
    flipped = bytearray(raw)
    at = 64 + ((len(text) + 1) * 8 + 63) // 64 * 64 + (len(text) * 4 + 63) // 64 * 64
    flipped[at] ^= 0xFF                                            # a bitmap byte of row 0: its payload no longer fits
    bad = str(tmp_path / "bits.psbrows")
    open(bad, "wb").write(bytes(flipped))
    for path, F in ((cut, 96), (bad, 96), (rows, 64)):             # ... and a file written for another width
        prod = native.StreamProducer(path, 1, 0, 1, F, 0, batch_rows=64, num_batches=3, packed="palette")
        with pytest.raises(native.NativeError, match="rows file"):
            _drain(prod, packed=True, timeout=10)
        prod.stop()


def test_converter_and_runner_take_rows_files(tmp_path):
    gen = ReviewGenerator(num_features=64, vocab=500, mean_len=15)
    write_csv(str(tmp_path / "train.csv"), gen.train_split(300))
    write_csv(str(tmp_path / "test.csv"), gen.test_split(100))
    env = dict(os.environ, PYTHONPATH=ROOT)
    for name in ("train", "test"):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "csv_to_rows.py"), str(tmp_path / f"{name}.csv"),
                            str(tmp_path / f"{name}.psbrows")], env=env, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0 and "B per row" in r.stdout, r.stderr
    logs = {}
    for ext in ("csv", "psbrows"):
        d = tmp_path / ext
        d.mkdir()
        r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner", "-training", str(tmp_path / f"train.{ext}"),
                            "-test", str(tmp_path / f"test.{ext}"), "--num_features", "64", "-c", "0", "-p", "0",
                            "--num_workers", "2", "--in_process", "--max_clock", "3", "-l", "-min", "8", "-max", "64"],
                           env=env, cwd=d, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        logs[ext] = parse_log(str(d / "logs-server.csv"), "server")
    assert len(logs["csv"]) == len(logs["psbrows"]) == 3
    assert np.allclose(logs["csv"][:, 4:], logs["psbrows"][:, 4:], atol=0.03)      # fp16 values vs the CSV's decimals
