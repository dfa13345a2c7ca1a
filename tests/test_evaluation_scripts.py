"""The evaluation side (the reference's logs and notebooks): the shipped logs carry the behaviour
the README reports, the same scripts read the reference's own logs unchanged, the plots come out headless."""
import importlib.util
import os
import struct
import subprocess
import sys
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOGS = os.path.join(ROOT, "evaluation", "logs")
REF_LOGS = "/root/reference/evaluation/logs"
RUNS = ["single-worker_5tps", "4-workers_0-5tps", "4-workers_2-5tps", "4-workers_5tps", "4-workers_10tps",
        "sequential", "bounded_delay_10", "eventual"]


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "evaluation", name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_shipped_logs_have_the_reference_run_names_and_the_reported_behaviour():
    sm = _load("summarize")
    rows = {r: sm.summarize(os.path.join(LOGS, r)) for r in RUNS}          # the reference's eight runs, same names
    assert rows["single-worker_5tps"]["workers"] == 1 and all(rows[r]["workers"] == 4 for r in RUNS[1:])
    # consistency models: clock gap 1 / k + 1 / unbounded (README.md:164-204 of the reference), F1 about the same
    assert rows["sequential"]["max clock gap"] == 1
    assert rows["bounded_delay_10"]["max clock gap"] == 11
    assert rows["eventual"]["max clock gap"] > 11
    f1 = [rows[r]["F1 mean last 50"] for r in ("sequential", "bounded_delay_10", "eventual")]
    assert max(f1) - min(f1) < 0.01 and min(f1) > 0.5
    # more events per second give a better model at the same wall-clock time
    by_rate = [rows[r]["F1 mean last 50"] for r in ("4-workers_0-5tps", "4-workers_2-5tps", "4-workers_10tps")]
    assert by_rate[0] < by_rate[1] < by_rate[2]
    table = sm.table(LOGS)
    assert table.count("\n") == 1 + len(RUNS) and "| sequential |" in table


@pytest.mark.skipif(not os.path.isdir(REF_LOGS), reason="reference checkout not present")
def test_the_same_scripts_read_the_reference_logs_unchanged():
    """Byte-compatible log formats work in both directions: the reference's published logs go through our reader
    (stray stdout lines such as the producer's EOF notice are skipped) and give BASELINE.md's numbers."""
    sm, pg = _load("summarize"), _load("plot_generation")
    seq = sm.summarize(os.path.join(REF_LOGS, "sequential"))
    assert seq["workers"] == 4 and seq["max clock gap"] == 1 and 0.35 < seq["F1 mean last 50"] < 0.45
    bd = sm.summarize(os.path.join(REF_LOGS, "bounded_delay_10"))
    ev = sm.summarize(os.path.join(REF_LOGS, "eventual"))
    assert 1 < bd["max clock gap"] <= 11 < ev["max clock gap"]
    run = pg.load_run(os.path.join(REF_LOGS, "sequential"))
    assert len(run["x"]) == seq["server clocks"] and np.nanmax(run["x"]) > 1000


def _png_size(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n" and data[12:16] == b"IHDR"
    w, h = struct.unpack(">II", data[16:24])
    pos, idat = 8, b""
    while pos < len(data):                                        # every chunk's CRC, and the pixels inflate
        n, kind = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(kind + body)
        idat += body if kind == b"IDAT" else b""
        pos += 12 + n
    assert len(zlib.decompress(idat)) >= w * h
    return w, h


def test_plots_are_written_headless(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "evaluation", "plot_generation.py"), LOGS, str(tmp_path)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    names = sorted(f for f in os.listdir(tmp_path) if f.endswith(".png"))
    assert names == sorted(f"{g}_{m}.png" for g in ("consistency", "event_rate", "num_workers")
                           for m in ("f1", "accuracy", "loss"))
    for n in names:
        w, h = _png_size(os.path.join(tmp_path, n))
        assert w >= 600 and h >= 360
    assert "consistency: sequential, bounded_delay_10, eventual" in open(tmp_path / "LEGEND.txt").read()


def test_ground_truth_trainer_reports_like_the_reference_notebook():
    """python-ground-truth-algorithm.ipynb prints one speed / validation line per epoch and a classification report;
    the offline trainer on the shipped small sample does the same."""
    sample = os.path.join(ROOT, "mockData", "small-sample")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "evaluation", "ground_truth.py"),
                        os.path.join(sample, "small_sample_reviews_embedded_training.csv"),
                        os.path.join(sample, "small_sample_reviews_embedded_test.csv")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    epochs = [ln for ln in r.stdout.splitlines() if ln.startswith("Epoch[")]
    assert len(epochs) == 10 and all("samples/sec" in ln and "Validation-Score-accuracy=" in ln for ln in epochs)
    assert "weighted avg" in r.stdout and "precision" in r.stdout
