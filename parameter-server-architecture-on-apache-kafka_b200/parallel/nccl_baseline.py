"""NCCL baseline (B0): the same worker math with library collectives instead of the fused protocol.

This is the measured baseline, not the product (BASELINE.json: "an NCCL-only path is the baseline").
Per clock: the worker kernel runs in ``PSB_COMM_NONE`` mode and writes one delta per worker, the deltas
of a GPU are summed with a PyTorch reduction, ``all_reduce`` (reduce-scatter + all-gather inside NCCL)
replaces push + pull, and ``psb_apply_delta`` performs ``w += delta / numWorkers``
(``ServerProcessor.java:148-151``).  Three kernels + one collective per clock instead of one kernel.
Only sequential consistency is expressible this way (a collective is a barrier).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import native
from ..engine.device_engine import DeviceEngine


class NcclBaselineEngine(DeviceEngine):
    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.w_dev = self.w0.clone()
        self.sum_delta = torch.zeros(self.nv, dtype=torch.float32, device=self.device)

    def connect(self) -> None:          # nothing to map: NCCL owns the transport
        return

    def step(self, mode=None) -> None:
        comm = native.comm_none()
        self.solver.inv_std = self._inv[self._win_i].data_ptr() if self.stats_at_ingest else None
        native.worker_step(self.solver, comm, self.W, self.impl, self.X, self.labels, self.win, self.w_dev, self.delta,
                           self.loss, self.info, self.scratch)
        torch.sum(self.delta, dim=0, out=self.sum_delta)
        if self.world > 1:
            dist.all_reduce(self.sum_delta)
        native.apply_delta(self.w_dev, self.sum_delta, self.lr, self.F, -1)
        self._step_done[self._steps & 1].record()         # (overlapped ring stores order themselves on these)
        self._steps += 1
        self._commits_since_step = 0
        self.kernel_launches += 2
        self.clock += 1

    def weights(self) -> torch.Tensor:
        return self.w_dev.view(self.layout.device_rows, 8)
