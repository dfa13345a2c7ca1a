"""BASELINE config 1: the shipped small sample through the in-process runner (sequential, 2 workers, CPU)."""
import os
import subprocess
import sys

import numpy as np

from psb200.utils.logs import parse_log

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SAMPLE = os.path.join(ROOT, "mockData", "small-sample")


def test_small_sample_sequential_two_workers(tmp_path):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner",
                        "-training", os.path.join(SAMPLE, "small_sample_reviews_embedded_training.csv"),
                        "-test", os.path.join(SAMPLE, "small_sample_reviews_embedded_test.csv"),
                        "-c", "0", "-p", "0", "--num_workers", "2", "--in_process", "--max_clock", "4", "-l",
                        "-min", "16", "-max", "64"], env=env, cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    s = parse_log(str(tmp_path / "logs-server.csv"), "server")
    w = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert len(s) == 4 and len(w) >= 6 and np.all(w[:, 6] == 60)
    assert np.all((s[:, 4] >= 0) & (s[:, 4] <= 1))


import pytest


def test_headerless_binary_label_file_trains_with_no_header(tmp_path):
    """The reference ships two more mock files (headerless, binary last column) that none of its code reads.  A file
    of that kind streams through the same runner with --no_header, labels 0 / 1 as class rows 0 / 1."""
    name, features = "headerless_binary_label.csv", 12
    path = os.path.join(ROOT, "mockData", name)
    rows = np.loadtxt(path, delimiter=",", ndmin=2)
    assert rows.shape == (400, features + 1) and set(np.unique(rows[:, -1])) == {0.0, 1.0}
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner", "-training", path, "-test", path,
                        "--no_header", "--num_features", str(features), "--num_classes", "1", "-c", "0", "-p", "0",
                        "--num_workers", "2", "--in_process", "--max_clock", "4", "-l", "-min", "8", "-max", "64"],
                       env=env, cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    s = parse_log(str(tmp_path / "logs-server.csv"), "server")
    w = parse_log(str(tmp_path / "logs-worker.csv"), "worker")
    assert len(s) == 4 and len(w) >= 6
    assert w[:, 6].max() == len(rows) // 2                  # every row of the file reached a worker (no line skipped)
    assert np.all(np.isfinite(s[:, 4])) and s[-1, 5] > 0.7  # a linear rule with noise: well above chance
