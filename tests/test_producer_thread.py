"""The C++ producer thread: partition filter, fp16 rows, batching, pacing, EOF (CsvProducer.java:36-87)."""
import os
import time

import numpy as np

from psb200.data import ReviewGenerator, read_csv, write_csv
from psb200.ops import native


def _drain(prod, timeout=20.0):
    rows, labels, workers, times = [], [], [], []
    t0 = time.time()
    while time.time() - t0 < timeout:
        b, n, eof = prod.poll()
        if b is None:
            if eof:
                break
            time.sleep(0.002)
            continue
        rows.append(prod.rows[b, :n].clone()), labels.append(prod.labels[b, :n].clone())
        workers.append(prod.workers[b, :n].clone()), times.append(prod.times[b, :n].clone())
        prod.release(b)
    cat = lambda xs: np.concatenate([x.numpy() for x in xs]) if xs else np.zeros(0)
    return cat(rows), cat(labels), cat(workers), cat(times), eof


def test_unthrottled_partition_slice(tmp_path):
    b = ReviewGenerator(num_features=64, vocab=500, mean_len=12).train_split(500)
    path = os.path.join(tmp_path, "t.csv")
    write_csv(path, b)
    prod = native.StreamProducer(path, 4, 1, 2, 64, 0, batch_rows=64, num_batches=3)   # partitions 1 and 2 of 4
    rows, labels, workers, times, eof = _drain(prod)
    prod.stop()
    assert eof == 1
    idx = np.nonzero(np.isin(np.arange(500) % 4, (1, 2)))[0]
    assert len(rows) == len(idx) == 250
    assert np.array_equal(labels, b.labels[idx]) and np.array_equal(workers, (idx % 4) - 1)
    assert np.allclose(rows.astype(np.float32), b.to_dense(np.float16)[idx].astype(np.float32), atol=2e-3)


def test_pacing_after_warmup(tmp_path):
    b = ReviewGenerator(num_features=16, vocab=100, mean_len=6).train_split(128 + 30)
    path = os.path.join(tmp_path, "t.csv")
    write_csv(path, b)
    prod = native.StreamProducer(path, 1, 0, 1, 16, 100, batch_rows=32, num_batches=4)   # 10 rows / s after 128 rows
    t0 = time.time()
    rows, labels, workers, times, eof = _drain(prod, timeout=15)
    took = time.time() - t0
    prod.stop()
    assert eof == 1 and len(rows) == 158
    assert times[127] < 500.0                        # warm-up rows arrive at once
    assert 2.5 < took < 12.0                         # 30 paced rows at 10 rows/s -> ~3 s (upper bound: a busy box)
    assert times[-1] - times[128] > 1900.0


def test_missing_file_reports_error(tmp_path):
    prod = native.StreamProducer(os.path.join(tmp_path, "nope.csv"), 1, 0, 1, 16, 0)
    _, _, _, _, eof = _drain(prod, timeout=5)
    prod.stop()
    assert eof == 2


import pytest


@pytest.mark.parametrize("mode", [True, "palette"])
def test_packed_batches_expand_to_the_dense_rows(tmp_path, mode):
    """StreamProducer(packed=True | "palette"): every published batch also carries its packed form, bit-identical
    to the dense fp16 rows of the same batch."""
    b = ReviewGenerator(num_features=64, vocab=500, mean_len=12).train_split(300)
    path = os.path.join(tmp_path, "t.csv")
    write_csv(path, b)
    prod = native.StreamProducer(path, 2, 0, 2, 64, 0, batch_rows=50, num_batches=3, packed=mode)
    assert prod.is_packed and prod.is_palette == (mode == "palette")
    seen, t0 = 0, time.time()
    while time.time() - t0 < 20:
        bi, n, eof = prod.poll()
        if bi is None:
            if eof:
                break
            time.sleep(0.002)
            continue
        pk = prod.packed(bi, n)
        dense = prod.rows[bi, :n].numpy()
        assert pk.n == n and (pk.palette or pk.nnz == int((dense != 0).sum()))
        assert np.array_equal(pk.to_dense().view(np.uint16), dense.view(np.uint16))
        assert pk.labels[:n].tolist() == prod.labels[bi, :n].tolist()
        seen += n
        prod.release(bi)
    prod.stop()
    assert seen == 300


@pytest.mark.parametrize("threads", [1, 3])
def test_tokenising_helper_threads_give_the_rows_of_the_single_thread(tmp_path, monkeypatch, threads):
    """PSB_PRODUCER_THREADS: the reading thread cuts the lines of a batch, helper threads tokenise them side by
    side, each line into its own row — rows, labels, partitions and their order must not depend on the thread
    count (CsvProducer.java:46-61 is one thread; the lines are independent).  The last line has no newline."""
    F, n = 96, 2500
    b = ReviewGenerator(num_features=F, vocab=700, mean_len=20).train_split(n)
    path = os.path.join(tmp_path, "t.csv")
    write_csv(path, b)
    with open(path, "rb+") as f:                     # cut the final newline off
        f.seek(-1, os.SEEK_END)
        if f.read(1) == b"\n":
            f.seek(-1, os.SEEK_END)
            f.truncate()
    monkeypatch.setenv("PSB_PRODUCER_THREADS", str(threads))
    prod = native.StreamProducer(path, 3, 0, 2, F, 0, batch_rows=256, num_batches=3, packed="palette")  # partitions 0, 1 of 3
    rows, labels, workers, times, eof = _drain(prod)
    prod.stop()
    idx = np.nonzero(np.arange(n) % 3 < 2)[0]
    assert eof == 1 and len(rows) == len(idx)
    assert np.array_equal(labels, b.labels[idx]) and np.array_equal(workers, idx % 3)
    text = read_csv(path).to_dense(np.float32)        # the values as written (6 decimals), rounded to fp16 once
    assert np.array_equal(rows.view(np.uint16), text[idx].astype(np.float16).view(np.uint16))
    assert (np.diff(times) >= 0).all()
