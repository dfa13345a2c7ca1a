"""The two CSV log streams, byte-compatible with the reference.

Server rows (``ServerProcessor.java:158-164``): ``<ms>;-1;<vc>;-1;<f1>;<accuracy>``
— partition and loss are always -1; one row per gradient received from worker 0 (Q6).
Worker rows (``WorkerTrainingProcessor.java:85-92``):
``<ms>;<partition>;<vc>;<loss>;<f1>;<accuracy>;<numTuplesSeen>``.
With ``-l`` stdout is redirected to ``./logs-server.csv`` / ``./logs-worker.csv`` and
the header is the first line (``ServerAppRunner.java:78-82``, ``WorkerAppRunner.java:77-81``);
everything printed afterwards lands in the file, including the producer's EOF line.
"""
from __future__ import annotations

import sys
import threading
from typing import List, Optional, TextIO



def _num(x) -> str:
    """Java's Double.toString for the values we print is close to Python's repr."""
    if isinstance(x, int):
        return str(x)
    return repr(float(x))


class LogSink:
    def __init__(self, stream: Optional[TextIO] = None, header: Optional[str] = None, keep: bool = False):
        self.stream = stream if stream is not None else sys.stdout
        self.lock = threading.Lock()
        self.rows: Optional[List[str]] = [] if keep else None
        self.observers: list = []                 # callables fed every line (utils/monitoring.py: --metrics_port)
        if header is not None:
            self.println(header)

    @staticmethod
    def to_file(path: str, header: str, keep: bool = False) -> "LogSink":
        return LogSink(open(path, "w", buffering=1), header, keep)

    def println(self, line: str) -> None:
        with self.lock:
            self.stream.write(line + "\n")
            self.stream.flush()
            if self.rows is not None:
                self.rows.append(line)
        for obs in self.observers:
            obs(line)

    def close(self) -> None:
        if self.stream not in (sys.stdout, sys.stderr):
            self.stream.close()


def server_row(ts_ms: int, vector_clock: int, f1: float, accuracy: float) -> str:
    return f"{int(ts_ms)};-1;{int(vector_clock)};-1;{_num(f1)};{_num(accuracy)}"


def worker_row(ts_ms: int, partition: int, vector_clock: int, loss: float, f1: float, accuracy: float,
               num_tuples_seen: int) -> str:
    return (f"{int(ts_ms)};{int(partition)};{int(vector_clock)};{_num(loss)};{_num(f1)};"
            f"{_num(accuracy)};{int(num_tuples_seen)}")


def parse_log(path_or_lines, kind: str):
    """Tolerant reader (skips the header and stray stdout lines such as the EOF notice)."""
    import numpy as np
    ncol = 6 if kind == "server" else 7
    lines = open(path_or_lines).read().splitlines() if isinstance(path_or_lines, str) else list(path_or_lines)
    rows = []
    for ln in lines:
        parts = ln.split(";")
        if len(parts) != ncol:
            continue
        try:
            rows.append([float(p) for p in parts])
        except ValueError:
            continue
    return np.asarray(rows, dtype=np.float64).reshape(-1, ncol)


__all__ = ["LogSink", "server_row", "worker_row", "parse_log", "SERVER_LOG_HEADER", "WORKER_LOG_HEADER"]
