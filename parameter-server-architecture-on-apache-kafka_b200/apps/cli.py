"""Argument parsing shared by the two runners.

Reference CLIs: ``ServerAppRunner.java:19-64`` and ``WorkerAppRunner.java:17-62``
(Apache commons-cli ``BasicParser``: short names are single-dash words such as
``-training``, long names double-dash).  argparse accepts the same spellings.
``-h`` prints usage and exits 0; a stray positional argument prints usage and
exits 2 (``ServerAppRunner.java:42-54``).

Flags that the reference hard-codes are added as long options only, so the
reference command lines keep working unchanged.
"""
from __future__ import annotations

import argparse
import sys
from typing import List, Optional

from ..parallel import RUNNER_TRANSPORTS
from ..config import (NUM_CLASSES_DEFAULT, NUM_FEATURES_DEFAULT, NUM_WORKERS_DEFAULT, PSConfig,
                      TEST_DATA_FILE_PATH_DEFAULT, TRAINING_DATA_FILE_PATH_DEFAULT)


class _Parser(argparse.ArgumentParser):
    def error(self, message):                    # commons-cli path: usage + exit code 2
        self.print_help(sys.stdout)
        self.exit(2)


def _common(p: argparse.ArgumentParser) -> None:
    p.add_argument("-test", "--test_data_file_path", default=TEST_DATA_FILE_PATH_DEFAULT, metavar="arg",
                   help="The path to an csv file that is used as test data (to compute statistics)")
    p.add_argument("-v", "--verbose", action="store_true", help="If enabled, prints the parameter that are used")
    p.add_argument("-h", "--help", action="store_true", help="Show list of possible parameter")
    p.add_argument("-r", "--remote", action="store_true",
                   help="If disabled, 'localhost' is used as the IP for the broker")
    # former compile-time constants / new capabilities
    p.add_argument("--num_workers", type=int, default=NUM_WORKERS_DEFAULT)
    p.add_argument("--num_features", type=int, default=NUM_FEATURES_DEFAULT)
    p.add_argument("--num_classes", type=int, default=NUM_CLASSES_DEFAULT)
    p.add_argument("--local_solver", choices=["lbfgs2", "sgd"], default="lbfgs2")
    p.add_argument("--init", choices=["zero", "random"], default="zero")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--backend", choices=["numpy", "cuda"], default="numpy",
                   help="math backend: CPU oracle or sm_100a kernels")
    p.add_argument("--transport", choices=list(RUNNER_TRANSPORTS), default="auto",
                   help="p2p: device peer-memory mailboxes between the two apps (default of --backend cuda); "
                        "loopback: host event loop (default of --backend numpy)")
    p.add_argument("--device", type=int, default=0, help="CUDA device of this app (p2p transport)")
    p.add_argument("--impl", choices=["tc", "simt"], default="tc", help="worker kernel flavour (p2p transport)")
    p.add_argument("--clocks_per_launch", type=int, default=1,
                   help="relaxed consistency models, p2p transport: clocks every worker runs per launch at its own pace")
    p.add_argument("--no_worker_metrics", action="store_true",
                   help="p2p transport: leave F1 / accuracy of the worker rows at -1 instead of scoring every local model")
    p.add_argument("--rendezvous", default=None, help="override the rendezvous address host:port")
    p.add_argument("--no_standardize", action="store_true")
    p.add_argument("--compat_q1", action="store_true", help="replicate the reference's off-by-one on key 6149")
    p.add_argument("--metrics_port", type=int, default=0,
                   help="serve Prometheus gauges of the log rows (clocks, loss, F1, accuracy, rows seen) on this port")
    p.add_argument("--no_header", action="store_true",
                   help="the CSV files have no header line (the reference always skips one, CsvProducer.java:41-43)")


def server_parser() -> argparse.ArgumentParser:
    p = _Parser(prog="ServerAppRunner", add_help=False)
    p.add_argument("-training", "--training_data_file_path", default=TRAINING_DATA_FILE_PATH_DEFAULT, metavar="arg",
                   help="The path to an csv file that is used as training data.")
    p.add_argument("-c", "--consistency_model", type=int, default=0, metavar="arg",
                   help="A number that defines the consistency model (see README for details)")
    p.add_argument("-p", "--producer_time_per_event", type=int, default=200, metavar="arg",
                   help="This is used to artificially increase/decrease the amount of events that created.")
    p.add_argument("-l", "--logging", action="store_true",
                   help="If enabled, writes performance logs into ./logs-server.csv")
    _common(p)
    p.add_argument("--in_process", action="store_true", help="host the workers in this process (no WorkerAppRunner)")
    p.add_argument("--max_clock", type=int, default=-1, help="stop after this many server clocks")
    p.add_argument("--free_running", action="store_true",
                   help="p2p transport: workers do not wait for the server's log row of a clock before the next one "
                        "(the server then logs the clocks it gets to see)")
    p.add_argument("--start_delay_ms", type=float, default=0.0,
                   help="delay before the training loop starts (reference: 20000)")
    p.add_argument("--virtual_clock", action="store_true", help="deterministic virtual time instead of wall clock")
    p.add_argument("--inject_delay", default="", help="straggler injection rank:ms[,rank:ms]")
    p.add_argument("-min", "--min_buffer_size", type=int, default=128, metavar="arg", help=argparse.SUPPRESS)
    p.add_argument("-max", "--max_buffer_size", type=int, default=1024, metavar="arg", help=argparse.SUPPRESS)
    p.add_argument("-bc", "--buffer_size_coefficient", type=float, default=0.3, metavar="arg", help=argparse.SUPPRESS)
    return p


def worker_parser() -> argparse.ArgumentParser:
    p = _Parser(prog="WorkerAppRunner", add_help=False)
    p.add_argument("-min", "--min_buffer_size", type=int, default=128, metavar="arg",
                   help="The minimum size for the buffer that stores the latest input events.")
    p.add_argument("-max", "--max_buffer_size", type=int, default=1024, metavar="arg",
                   help="The maximum size for the buffer that stores the latest input events.")
    p.add_argument("-bc", "--buffer_size_coefficient", type=float, default=0.3, metavar="arg",
                   help="buffer_size = coefficient * events_per_minute")
    p.add_argument("-l", "--logging", action="store_true",
                   help="If enabled, writes performance logs into ./logs-worker.csv")
    _common(p)
    return p


def config_from_args(ns: argparse.Namespace, role: str) -> PSConfig:
    cfg = PSConfig(num_workers=ns.num_workers, num_features=ns.num_features, num_classes=ns.num_classes,
                   test_data_file_path=ns.test_data_file_path, local_solver=ns.local_solver, init=ns.init,
                   seed=ns.seed, transport=ns.transport, remote=ns.remote, verbose=ns.verbose,
                   logging=ns.logging, standardize=not ns.no_standardize, compat_q1=ns.compat_q1, csv_header=not ns.no_header,
                   min_buffer_size=ns.min_buffer_size, max_buffer_size=ns.max_buffer_size,
                   buffer_size_coefficient=ns.buffer_size_coefficient)
    if role == "server":
        cfg = cfg.replace(training_data_file_path=ns.training_data_file_path,
                          consistency_model=ns.consistency_model,
                          producer_time_per_event=ns.producer_time_per_event, max_clock=ns.max_clock,
                          inject_delay=ns.inject_delay)
    return cfg


def parse(parser: argparse.ArgumentParser, argv: Optional[List[str]]):
    ns = parser.parse_args(argv)
    if ns.help:
        parser.print_help(sys.stdout)
        raise SystemExit(0)
    if getattr(ns, "transport", None) == "auto":
        ns.transport = "p2p" if getattr(ns, "backend", "numpy") == "cuda" else "loopback"
    if getattr(ns, "transport", None) == "p2p" and getattr(ns, "backend", "cuda") != "cuda":
        parser.error("--transport p2p needs --backend cuda")
    return ns
