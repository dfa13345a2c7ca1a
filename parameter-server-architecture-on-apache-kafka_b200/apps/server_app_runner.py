"""``ServerAppRunner`` entry point — ``python -m psb200.apps.server_app_runner``.

Reference: ``ServerAppRunner.java:14-102``.  Same options and defaults, same
verbose block (``:67-76``), same ``-l`` redirection of stdout to ``./logs-server.csv``
with the schema line first (``:78-82``).  The reference's 20 s sleep before the
topology starts (``:95``) is ``--start_delay_ms`` (default 0: the rendezvous makes
it unnecessary).
"""
from __future__ import annotations

import sys
from typing import List, Optional

from ..config import RENDEZVOUS_LOCAL, RENDEZVOUS_REMOTE, SERVER_LOG_HEADER, WORKER_LOG_HEADER
from ..utils.logs import LogSink
from . import cli
from .server_app import ServerApp


def main(argv: Optional[List[str]] = None) -> int:
    ns = cli.parse(cli.server_parser(), argv)
    from ..utils.logging_config import configure
    configure()
    cfg = cli.config_from_args(ns, "server")
    broker = ns.rendezvous or (RENDEZVOUS_REMOTE if cfg.remote else RENDEZVOUS_LOCAL)
    if cfg.verbose:
        print()
        print("Used parameter:")
        print(f"    training_data_file_path: {cfg.training_data_file_path}")
        print(f"    test_data_file_path: {cfg.test_data_file_path}")
        print(f"    consistency_model: {cfg.consistency_model}")
        print(f"    producer_time_per_event: {cfg.producer_time_per_event}")
        print(f"    broker address: {broker}")
        print()
    server_log = LogSink.to_file("./logs-server.csv", SERVER_LOG_HEADER) if cfg.logging else LogSink()
    worker_log = None
    if ns.in_process:
        worker_log = LogSink.to_file("./logs-worker.csv", WORKER_LOG_HEADER) if cfg.logging else LogSink()
    from ..utils.monitoring import start_exporter
    exporter = start_exporter(ns.metrics_port, server_log, worker_log)
    if ns.rendezvous:
        import psb200.config as c
        c.RENDEZVOUS_LOCAL = c.RENDEZVOUS_REMOTE = ns.rendezvous
    app = ServerApp(cfg, in_process=ns.in_process, backend=ns.backend, server_log=server_log,
                    worker_log=worker_log, virtual=ns.virtual_clock, start_delay_ms=ns.start_delay_ms, ns=ns)
    app.call()
    server_log.close()
    if worker_log is not None:
        worker_log.close()
    if exporter is not None:
        exporter.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
