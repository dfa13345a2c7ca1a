"""CUDA math backend for the event-driven engine (``PSSimulation``) and the two Runner CLIs.

Same interface as ``NumpyBackend``; every worker step is one ``psb_worker_step`` launch in
``PSB_COMM_NONE`` mode (the engine applies the delta, as ``ServerProcessor.java:148-151``
does), evaluation is ``psb_eval``.  Used for reference-rate streaming runs where the
scheduler lives on the host; the throughput path is ``DeviceEngine``.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from ..config import PSConfig
from ..data.csr import CSRBatch
from ..models.logreg import LogRegLayout
from ..ops import native
from ..utils.metrics import Metrics


class CudaBackend:
    name = "cuda"

    def __init__(self, cfg: PSConfig, test: Optional[CSRBatch], impl: int = native.IMPL_SIMT):
        if not torch.cuda.is_available():
            raise native.NativeError("--backend cuda needs a CUDA device")
        native.lib()
        self.cfg, self.impl = cfg, impl
        self.layout = LogRegLayout(cfg.num_features, cfg.num_classes)
        dev = torch.device("cuda")
        F, cap, W = cfg.num_features, cfg.max_buffer_size, cfg.num_workers
        self.F, self.cap, self.nv = F, cap, self.layout.device_size
        self.X = torch.zeros((W, cap, F), dtype=torch.float16, device=dev)
        self.labels = torch.zeros((W, cap), dtype=torch.int32, device=dev)
        self.win = torch.zeros((W, 2), dtype=torch.int32, device=dev)
        self.w_dev = torch.zeros(self.nv, dtype=torch.float32, device=dev)
        self.w_in = torch.zeros(self.nv, dtype=torch.float32, device=dev)
        self.delta = torch.zeros((1, self.nv), dtype=torch.float32, device=dev)
        self.loss = torch.zeros(1, dtype=torch.float32, device=dev)
        self.info = torch.zeros((1, 8), dtype=torch.int32, device=dev)
        self.scratch = torch.zeros(native.scratch_floats(F), dtype=torch.float32, device=dev)
        self.conf = torch.zeros(64, dtype=torch.int32, device=dev)
        self.mout = torch.zeros(2, dtype=torch.float32, device=dev)
        self.row = torch.zeros(F, dtype=torch.float32, device=dev)
        self.solver = native.make_solver(F, self.layout.R, cap, cfg.local_solver, cfg.max_iter, cfg.standardize,
                                         cfg.center, cfg.sgd_lr)
        self.weights = self.layout.zeros() if cfg.init == "zero" else self.layout.random(cfg.seed)
        self.w_dev.copy_(torch.from_numpy(self.weights.reshape(-1)))
        if test is not None:
            self.test_X = torch.from_numpy(test.to_dense(np.float16)).to(dev)
            self.test_y = torch.from_numpy(test.labels.astype(np.int32)).to(dev)
        else:
            self.test_X = None
        self.last_info = None

    STAGE_ROWS = 256

    def store_row(self, worker: int, slot: int, idx: np.ndarray, val: np.ndarray, label: int) -> None:
        """Rows are collected in pinned staging memory and stored into the rings by ONE copy + ONE store kernel
        (``psb_ring_store_dense``) when a worker step needs them or the staging set is full."""
        if not hasattr(self, "_stage"):
            self._stage = torch.zeros((self.STAGE_ROWS, self.F), dtype=torch.float16).pin_memory()
            self._stage_lab = torch.zeros(self.STAGE_ROWS, dtype=torch.int32).pin_memory()
            self._stage_ws = torch.zeros((self.STAGE_ROWS, 2), dtype=torch.int32).pin_memory()
            dev = self.X.device
            self._stage_dev = torch.zeros((self.STAGE_ROWS, self.F), dtype=torch.float16, device=dev)
            self._stage_lab_dev = torch.zeros(self.STAGE_ROWS, dtype=torch.int32, device=dev)
            self._stage_ws_dev = torch.zeros((self.STAGE_ROWS, 2), dtype=torch.int32, device=dev)
            self._staged = 0
        k = self._staged
        row = self._stage[k].numpy()
        row[:] = 0
        row[idx] = val
        self._stage_lab[k] = int(label)
        self._stage_ws[k, 0], self._stage_ws[k, 1] = int(worker), int(slot)
        # an earlier staged row of the same slot is superseded (the store kernel writes the rows of a batch in parallel)
        ws = self._stage_ws[:k].numpy()
        if k:
            ws[(ws[:, 0] == worker) & (ws[:, 1] == slot), 1] = -1
        self._staged = k + 1
        if self._staged == self.STAGE_ROWS:
            self.flush_rows()

    def flush_rows(self) -> None:
        n = getattr(self, "_staged", 0)
        if not n:
            return
        self._stage_dev[:n].copy_(self._stage[:n], non_blocking=True)
        self._stage_lab_dev[:n].copy_(self._stage_lab[:n], non_blocking=True)
        self._stage_ws_dev[:n].copy_(self._stage_ws[:n], non_blocking=True)
        native.ring_store_dense(self._stage_dev, self._stage_lab_dev, self._stage_ws_dev, n, self.X, self.labels,
                                self.cap, self.F)
        torch.cuda.current_stream().synchronize()            # the pinned staging set is reused right away
        self._staged = 0

    def snapshot(self) -> np.ndarray:
        return self.w_dev.cpu().numpy().reshape(self.layout.device_rows, 8).copy()

    def _metrics(self, w: torch.Tensor) -> Optional[Metrics]:
        if self.test_X is None:
            return None
        native.evaluate(self.test_X, self.test_y, self.test_X.shape[0], w, self.F, self.layout.R, self.conf, self.mout,
                        self.impl)
        f1, acc = self.mout.cpu().tolist()
        return Metrics(f1, acc)

    def worker_step(self, worker: int, weights: np.ndarray, slots: List[int], want_metrics: bool):
        self.flush_rows()
        head = (slots[-1] + 1) % self.cap
        self.win[worker, 0], self.win[worker, 1] = head, len(slots)
        self.w_in.copy_(torch.from_numpy(np.ascontiguousarray(weights, dtype=np.float32).reshape(-1)))
        comm = native.comm_none()
        native.worker_step(self.solver, comm, 1, self.impl, self.X[worker], self.labels[worker], self.win[worker],
                           self.w_in, self.delta, self.loss, self.info, self.scratch)
        delta = self.delta[0].cpu().numpy().reshape(self.layout.device_rows, 8).copy()
        loss = float(self.loss.item())
        self.last_info = self.info[0].cpu().tolist()
        m = self._metrics(self.w_in + self.delta[0]) if want_metrics else None
        return delta, loss, m

    def apply(self, delta: np.ndarray, lr: float) -> None:
        d = torch.from_numpy(np.ascontiguousarray(delta, dtype=np.float32).reshape(-1)).to(self.w_dev.device)
        skip = self.layout.R - 1 if self.cfg.compat_q1 else -1
        native.apply_delta(self.w_dev, d, float(lr), self.F, skip)

    def server_metrics(self) -> Optional[Metrics]:
        return self._metrics(self.w_dev)
