"""Checkpoint round trip, the progress watchdog and the metrics endpoint (CPU)."""
import os
import time

import numpy as np

from psb200.models import LogRegLayout
from psb200.utils import checkpoint
from psb200.utils.watchdog import Watchdog


def test_checkpoint_roundtrip_uses_reference_keys(tmp_path):
    lay = LogRegLayout(64, 5)
    W = lay.random(3)
    path = checkpoint.save(str(tmp_path), 7, W, lay, clocks=[7, 7], rows_ingested=123)
    lines = open(path).read().splitlines()
    assert lines[0] == "key;value" and len(lines) == 1 + lay.num_params
    k, v = lines[1 + lay.key_of(3, 2)].split(";")
    assert int(k) == lay.key_of(3, 2) and abs(float(v) - W[3, 2]) < 1e-7
    W2, meta = checkpoint.load(str(tmp_path), lay)
    assert np.array_equal(W2, W) and meta["clock"] == 7 and meta["rows_ingested"] == 123


def test_watchdog_trips_on_stall_and_on_status():
    msgs = []
    wd = Watchdog(lambda: 5, lambda: 0, timeout_s=0.3, on_stall=msgs.append, poll_s=0.05).start()
    time.sleep(0.8)
    wd.stop()
    assert wd.tripped and "no progress" in msgs[0]
    msgs.clear()
    wd = Watchdog(lambda: int(time.time() * 100), lambda: 2, timeout_s=5, on_stall=msgs.append, poll_s=0.05).start()
    time.sleep(0.3)
    wd.stop()
    assert wd.tripped and "status word 2" in msgs[0]


def test_watchdog_quiet_while_progressing():
    wd = Watchdog(lambda: int(time.time() * 100), lambda: 0, timeout_s=0.2, poll_s=0.05).start()
    time.sleep(0.6)
    wd.stop()
    assert wd.tripped is None


def test_metrics_endpoint_follows_the_log_rows(tmp_path):
    """--metrics_port: a Prometheus endpoint fed by the same rows the CSV logs get (the reference watches message
    flow from outside through monitoring interceptors, BaseKafkaApp.java:73-78).  Runs the in-process server app in
    this process and scrapes its endpoint afterwards."""
    import socket
    import urllib.request

    from psb200.apps import server_app_runner
    from psb200.utils.logs import LogSink, server_row, worker_row
    from psb200.utils.monitoring import MetricsExporter

    with socket.socket() as s:                       # a free port
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    exp = MetricsExporter(port, addr="127.0.0.1")
    sink = LogSink(stream=open(os.devnull, "w"))
    exp.attach(sink, None)
    sink.println("timestamp;partition;vectorClock;loss;fMeasure;accuracy;numTuplesSeen")      # header: ignored
    sink.println(worker_row(1, 0, 7, 0.5, 0.41, 0.45, 300))
    sink.println(worker_row(2, 1, 4, 0.75, 0.40, 0.44, 290))
    sink.println(server_row(3, 7, 0.43, 0.47))
    sink.println("CSV Producer sent all records!")
    text = urllib.request.urlopen(f"http://127.0.0.1:{port}/metrics", timeout=10).read().decode()
    exp.close()
    got = {ln.split(" ")[0]: float(ln.split(" ")[1]) for ln in text.splitlines() if ln and not ln.startswith("#")}
    assert got["psb_server_clock"] == 7 and got["psb_server_f1"] == 0.43 and got["psb_server_rows_total"] == 1
    assert got['psb_worker_clock{partition="1"}'] == 4 and got['psb_worker_loss{partition="1"}'] == 0.75
    assert got['psb_worker_tuples_seen{partition="0"}'] == 300 and got["psb_worker_rows_total"] == 2
    assert got['psb_clock_lag{partition="1"}'] == 3 and got['psb_clock_lag{partition="0"}'] == 0

    # the flag on the runner: the endpoint lives as long as the app; rows flow through it without changing the logs
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        rc = server_app_runner.main(["-training", "synthetic:200", "-test", "synthetic:80", "-p", "0", "--num_workers", "2",
                                     "--in_process", "--max_clock", "2", "-l", "--num_features", "64", "-min", "8",
                                     "-max", "32", "--metrics_port", str(port)])
    finally:
        os.chdir(cwd)
    from psb200.utils.logs import parse_log
    assert rc == 0 and len(parse_log(str(tmp_path / "logs-server.csv"), "server")) == 2
