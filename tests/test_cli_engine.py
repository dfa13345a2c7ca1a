"""Runner CLIs (options, defaults, exit codes, log headers) and the in-process / two-process engines."""
import os
import subprocess
import sys

import numpy as np
import pytest

from psb200.apps import cli
from psb200.config import PSConfig, SERVER_LOG_HEADER, WORKER_LOG_HEADER
from psb200.consistency import max_gap_bound
from psb200.data import ReviewGenerator
from psb200.engine.sim import PSSimulation
from psb200.utils.logs import parse_log

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))


def test_server_defaults_match_reference():
    ns = cli.server_parser().parse_args([])
    assert ns.training_data_file_path == "./data/train.csv" and ns.test_data_file_path == "./data/test.csv"
    assert ns.consistency_model == 0 and ns.producer_time_per_event == 200
    assert not (ns.verbose or ns.remote or ns.logging or ns.help)
    ns = cli.server_parser().parse_args(["-training", "a.csv", "-test", "b.csv", "-c", "-1", "-p", "50", "-v", "-r", "-l"])
    cfg = cli.config_from_args(ns, "server")
    assert (cfg.training_data_file_path, cfg.test_data_file_path, cfg.consistency_model,
            cfg.producer_time_per_event) == ("a.csv", "b.csv", -1, 50)
    assert cfg.remote and cfg.rendezvous == "kafka:9092" and cfg.logging and cfg.verbose
    ns = cli.server_parser().parse_args(["--training_data_file_path", "x", "--consistency_model", "10",
                                         "--producer_time_per_event", "7"])
    assert ns.training_data_file_path == "x" and ns.consistency_model == 10 and ns.producer_time_per_event == 7


def test_worker_defaults_match_reference():
    ns = cli.worker_parser().parse_args([])
    assert (ns.min_buffer_size, ns.max_buffer_size, ns.buffer_size_coefficient) == (128, 1024, 0.3)
    ns = cli.worker_parser().parse_args(["-min", "64", "-max", "256", "-bc", "0.5", "-test", "t.csv"])
    cfg = cli.config_from_args(ns, "worker")
    assert (cfg.min_buffer_size, cfg.max_buffer_size, cfg.buffer_size_coefficient) == (64, 256, 0.5)
    assert PSConfig().rendezvous == "127.0.0.1:29092"


def test_help_exits_0_and_stray_args_exit_2():
    for mod in ("server_app_runner", "worker_app_runner"):
        r = subprocess.run([sys.executable, "-m", f"psb200.apps.{mod}", "-h"], env=ENV, capture_output=True, text=True,
                           timeout=120)
        assert r.returncode == 0 and "usage:" in r.stdout
        r = subprocess.run([sys.executable, "-m", f"psb200.apps.{mod}", "stray"], env=ENV, capture_output=True,
                           text=True, timeout=120)
        assert r.returncode == 2 and "usage:" in r.stdout


def test_in_process_cli_writes_reference_log_formats(tmp_path):
    r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner", "-training", "synthetic:600", "-test",
                        "synthetic:300", "-p", "0", "--num_workers", "2", "--in_process", "--max_clock", "3", "-l",
                        "-v"], env=ENV, cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Used parameter:" in r.stdout and "consistency_model: 0" in r.stdout
    s = open(tmp_path / "logs-server.csv").read().splitlines()
    w = open(tmp_path / "logs-worker.csv").read().splitlines()
    assert s[0] == SERVER_LOG_HEADER and w[0] == WORKER_LOG_HEADER
    assert "CSV Producer sent all records!" in s
    srows, wrows = parse_log(s, "server"), parse_log(w, "worker")
    assert len(srows) == 3 and np.all(srows[:, 1] == -1) and np.all(srows[:, 3] == -1)
    assert list(srows[:, 2]) == [0, 1, 2]
    assert wrows.shape[1] == 7 and set(wrows[:, 1]) == {0, 1} and np.all(wrows[:, 6] == 300)


def test_two_process_gloo_matches_in_process(tmp_path):
    """BASELINE config 1: sequential, numWorkers=2, CPU/gloo through both Runner CLIs."""
    port = 29600 + os.getpid() % 300
    a, b = tmp_path / "two", tmp_path / "one"
    a.mkdir(), b.mkdir()
    common = ["-test", "synthetic:300", "--num_workers", "2"]
    wp = subprocess.Popen([sys.executable, "-m", "psb200.apps.worker_app_runner", *common, "-l", "--rendezvous",
                           f"127.0.0.1:{port}"], env=ENV, cwd=a)
    try:
        r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner", "-training", "synthetic:600",
                            *common, "-p", "0", "--max_clock", "4", "-l", "--rendezvous", f"127.0.0.1:{port}"],
                           env=ENV, cwd=a, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        assert wp.wait(timeout=60) == 0
    finally:
        if wp.poll() is None:
            wp.kill()
    r = subprocess.run([sys.executable, "-m", "psb200.apps.server_app_runner", "-training", "synthetic:600", *common,
                        "-p", "0", "--max_clock", "4", "-l", "--in_process"], env=ENV, cwd=b, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    s2, s1 = parse_log(str(a / "logs-server.csv"), "server"), parse_log(str(b / "logs-server.csv"), "server")
    assert len(s2) == 4 and np.allclose(s2[:, 2:], s1[:, 2:])           # same clocks, F1 and accuracy
    assert s2[-1, 4] > s2[0, 4]                                         # F1 rises
    w2 = parse_log(str(a / "logs-worker.csv"), "worker")
    assert len(w2) >= 8


@pytest.mark.parametrize("model", [0, 2, -1])
def test_straggler_injection_respects_consistency_model(model):
    g = ReviewGenerator(num_features=64, vocab=400, mean_len=10)
    tr, te = g.train_split(800), g.test_split(200)
    cfg = PSConfig(num_workers=3, num_features=64, consistency_model=model, producer_time_per_event=0, max_clock=15,
                   min_buffer_size=16, max_buffer_size=64, inject_delay="2:9000")
    sim = PSSimulation(cfg, tr, te, iter_time_ms=lambda w, c: 1000.0, server_start_ms=0.0)
    sim.run()
    bound = max_gap_bound(model)
    if bound is None:
        assert sim.max_gap > 3                          # eventual: the fast workers are unaffected
    else:
        assert sim.max_gap <= bound
    assert sim.server_clock == 15                       # no deadlock in any model


def test_sequential_is_deterministic_and_learns():
    g = ReviewGenerator()
    tr, te = g.train_split(2000), g.test_split(600)
    cfg = PSConfig(num_workers=2, producer_time_per_event=0, max_clock=8)
    runs = []
    for _ in range(2):
        sim = PSSimulation(cfg, tr, te, server_start_ms=0.0)
        sim.run()
        runs.append(parse_log(sim.server_log.rows, "server"))
    assert np.array_equal(runs[0], runs[1])
    assert runs[0][-1, 4] > runs[0][0, 4] and runs[0][-1, 4] > 0.3
