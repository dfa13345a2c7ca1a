"""Shared runtime base of the two apps: config, rendezvous, data loading, shutdown.

Reference: ``BaseKafkaApp.java:42-61`` (``call()``: build the streams runtime,
install a SIGINT shutdown hook, clean up, start) and ``:63-81`` (client
properties).  Here ``call()`` installs the same SIGINT hook, joins the
rendezvous when the app runs as one half of a two-process deployment and hands
over to the subclass' ``run``.
"""
from __future__ import annotations

import os
import signal

from ..config import PSConfig
from ..data import CSRBatch, ReviewGenerator, read_csv


def load_csv_or_synthetic(path: str, cfg: PSConfig, kind: str) -> CSRBatch:
    """Reads a dataset CSV; ``synthetic:<rows>[:<seed>]`` generates the stand-in instead."""
    if path.startswith("synthetic:"):
        parts = path.split(":")
        n = int(parts[1])
        seed = int(parts[2]) if len(parts) > 2 else cfg.seed
        gen = ReviewGenerator(cfg.num_features, cfg.num_classes, seed=seed)
        return gen.test_split(n, seed=seed + 2) if kind == "test" else gen.train_split(n, seed=seed + 1)
    if not os.path.exists(path):
        raise FileNotFoundError(f"{kind} data file not found: {path}")
    from ..data import is_rows_file, read_rows_file
    if is_rows_file(path):                     # a CSV converted once by tools/csv_to_rows.py
        return read_rows_file(path)
    try:
        from ..ops import native
        if native.available():
            return native.csv_to_csr(path, has_header=cfg.csv_header)
    except Exception:
        pass
    return read_csv(path, has_header=cfg.csv_header)


class BaseApp:
    application_id = "parameter-server"

    def __init__(self, cfg: PSConfig):
        self.cfg = cfg
        self._closing = False

    def run(self) -> None:
        raise NotImplementedError

    on_close = None                               # hook of the running transport (p2p: raise its stop flag)

    def close(self) -> None:
        self._closing = True
        if self.on_close is not None:
            self.on_close()

    def call(self) -> None:
        def _hook(signum, frame):
            self.close()
            raise KeyboardInterrupt
        try:
            signal.signal(signal.SIGINT, _hook)
        except ValueError:
            pass                                  # not the main thread (tests)
        try:
            self.run()
        except KeyboardInterrupt:
            pass
        finally:
            from ..parallel.link import close_link
            close_link()
