"""Worker app: sampling processor + training processor for all partitions.

Reference: ``WorkerApp.java:33-42`` wires ``INPUT_DATA -> WorkerSamplingProcessor``
and ``WEIGHTS_TOPIC -> WorkerTrainingProcessor`` around one state store.  Here the
app serves ROW / WEIGHTS / STOP messages from the server app over the link: a ROW
goes through the dynamic buffer (``WorkerSamplingProcessor.java:50-113``), a
WEIGHTS message triggers one local solve on that partition's buffer and a
GRADIENT reply plus one worker log row (``WorkerTrainingProcessor.java:63-98``).
"""
from __future__ import annotations

import time
from typing import Optional

import numpy as np

from ..config import PSConfig
from ..engine.buffer import DynamicBuffer
from ..utils.logs import LogSink, worker_row
from .base_app import BaseApp, load_csv_or_synthetic
from .server_app import make_backend


class WorkerApp(BaseApp):
    application_id = "parameter-server-worker"        # WorkerApp.java:46-48

    def __init__(self, cfg: PSConfig, backend: str = "numpy", worker_log: Optional[LogSink] = None, ns=None):
        super().__init__(cfg)
        self.backend_name = backend
        self.worker_log = worker_log or LogSink()
        self.ns = ns

    def run(self) -> None:
        if self.cfg.transport == "p2p":               # product path (psb200.apps.p2p_apps.P2PWorker)
            from .p2p_apps import run_worker
            run_worker(self, self.cfg, self.ns, self.worker_log)
            return
        from ..parallel.link import GRADIENT, ROW, STOP, WEIGHTS, WORKER_RANK, WorkerLink, init_link
        cfg = self.cfg
        test = load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test")
        backend = make_backend(cfg, test, self.backend_name)
        buffers = [DynamicBuffer(cfg.min_buffer_size, cfg.max_buffer_size, cfg.buffer_size_coefficient)
                   for _ in range(cfg.num_workers)]
        init_link(WORKER_RANK, cfg.rendezvous)
        link = WorkerLink()
        lay = backend.layout
        stop = False
        while not (stop or self._closing):
            replies = []
            for kind, worker, clock, ints, floats, aux in link.recv_batch():
                if kind == STOP:
                    stop = True
                elif kind == ROW:
                    slot = buffers[worker].on_event(time.time() * 1000.0)
                    backend.store_row(worker, slot, np.asarray(ints, dtype=np.int32),
                                      np.asarray(floats, dtype=np.float32), int(aux))
                elif kind == WEIGHTS:
                    slots = buffers[worker].ring.active_slots()
                    if not slots:
                        raise RuntimeError(f"There is no data for partition {worker}")
                    weights = np.asarray(floats, dtype=np.float32).reshape(lay.device_rows, lay.CP)
                    delta, loss, m = backend.worker_step(worker, weights, slots, True)
                    f1, acc = (m.f1, m.accuracy) if m is not None else (-1.0, -1.0)
                    seen = buffers[worker].ring.total
                    self.worker_log.println(worker_row(int(time.time() * 1000), worker, clock, loss, f1, acc, seen))
                    payload = np.concatenate([delta.reshape(-1), np.asarray([loss, f1, acc], dtype=np.float32)])
                    replies.append((GRADIENT, worker, clock, None, payload, seen))
                else:
                    raise RuntimeError(f"worker app received unexpected message kind {kind}")
            link.reply(replies)
