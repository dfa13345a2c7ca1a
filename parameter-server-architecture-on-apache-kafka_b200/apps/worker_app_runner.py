"""``WorkerAppRunner`` entry point — ``python -m psb200.apps.worker_app_runner``.

Reference: ``WorkerAppRunner.java:13-94``: options ``-test -min -max -bc -v -h -r -l``,
verbose block, ``-l`` redirection to ``./logs-worker.csv`` with the schema line
(``:77-81``), 10 s sleep (``:84``, not needed here: the rendezvous blocks until the
server app is up).  The help text of the reference mistakenly says
"ServerAppRunner" (``:43``, ``:50``); ours says WorkerAppRunner.
"""
from __future__ import annotations

import sys
from typing import List, Optional

from ..config import RENDEZVOUS_LOCAL, RENDEZVOUS_REMOTE, WORKER_LOG_HEADER
from ..utils.logs import LogSink
from . import cli
from .worker_app import WorkerApp


def main(argv: Optional[List[str]] = None) -> int:
    ns = cli.parse(cli.worker_parser(), argv)
    from ..utils.logging_config import configure
    configure()
    cfg = cli.config_from_args(ns, "worker")
    broker = ns.rendezvous or (RENDEZVOUS_REMOTE if cfg.remote else RENDEZVOUS_LOCAL)
    if cfg.verbose:
        print()
        print("Used parameter:")
        print(f"    test_data_file_path: {cfg.test_data_file_path}")
        print(f"    min_buffer_size: {cfg.min_buffer_size}")
        print(f"    max_buffer_size: {cfg.max_buffer_size}")
        print(f"    buffer_size_coefficient: {cfg.buffer_size_coefficient}")
        print(f"    broker address: {broker}")
        print()
    worker_log = LogSink.to_file("./logs-worker.csv", WORKER_LOG_HEADER) if cfg.logging else LogSink()
    from ..utils.monitoring import start_exporter
    exporter = start_exporter(ns.metrics_port, worker_log)
    if ns.rendezvous:
        import psb200.config as c
        c.RENDEZVOUS_LOCAL = c.RENDEZVOUS_REMOTE = ns.rendezvous
    app = WorkerApp(cfg, backend=ns.backend, worker_log=worker_log, ns=ns)
    app.call()
    worker_log.close()
    if exporter is not None:
        exporter.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
