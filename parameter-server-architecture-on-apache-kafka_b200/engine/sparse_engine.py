"""Sharded sparse engine (BASELINE config 5): 2^24 hashed features, server shards across the GPUs.

The dense engine replicates the 6150-parameter model on every GPU; here the model (F x 8 fp32, 512 MiB
at F = 2^24) is range-partitioned — rank s owns rows [s*Fs, (s+1)*Fs) — which is the key-range sharding
the reference designs for (``KeyRange`` in every message, ``BaseMessage.java:24-27``) but never turns on
(``ServerApp.java:38``, ``README.md:115-119``, future work ``README.md:333``).  A worker step is one SGD
step on a CSR batch (``psb_sparse_step``): the pull is the forward's operand loads from the owning shards,
the push is a 16-byte RED into them.  Sequential consistency separates readers from writers with a
device-side phase barrier over peer memory; eventual consistency fuses both into one kernel; bounded delay k
is that one kernel behind the clock gate ``min over ranks of completed steps >= v - k``
(``MessageTracker.java:69-79``), evaluated in the kernel on the ranks' step counters in peer memory —
independent of the model size, as in the reference (``ServerProcessor.java:126-131``).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from ..data.csr import CSRBatch
from ..ops import native
from .device_engine import _view


class DeviceCSR:
    def __init__(self, batch: CSRBatch, device):
        self.indptr = torch.from_numpy(batch.indptr.astype(np.int64)).to(device)
        self.indices = torch.from_numpy(batch.indices.astype(np.int32)).to(device)
        self.values = torch.from_numpy(batch.values.astype(np.float32)).to(device)
        self.labels = torch.from_numpy(batch.labels.astype(np.int32)).to(device)
        self.n, self.nnz = len(batch), batch.nnz


class SparseShardedEngine:
    def __init__(self, num_features: int, num_classes: int = 5, rank: int = 0, world: int = 1,
                 num_workers_total: Optional[int] = None, eta: float = 0.5, consistency_model: int = 0,
                 device: Optional[torch.device] = None, spin_limit_cycles: int = 20_000_000_000):
        if not torch.cuda.is_available():
            raise native.NativeError("SparseShardedEngine needs a CUDA device")
        native.lib()
        self.F, self.R = num_features, num_classes + 1
        self.rank, self.world = rank, world
        self.num_workers = num_workers_total or world
        self.eta, self.model = eta, consistency_model
        self.host_delay_s = 0.0                # straggler injection (tests, tools/straggler_sweep.py)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.Fs = (num_features + world - 1) // world
        self.shard_bytes = (self.Fs + 1) * 32
        self.shard = native.device_malloc(self.shard_bytes)
        self.ctrl = native.device_malloc(512)                    # [0..32) phases, [32] ticket, [33] status
        _view(self.shard, (self.Fs + 1) * 8, self.device).zero_()
        _view(self.ctrl, 128, self.device, torch.int32).zero_()
        self.shards: List[int] = [0] * world
        self.ctrls: List[int] = [0] * world
        self.shards[rank], self.ctrls[rank] = self.shard, self.ctrl
        self.loss = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.residuals: Optional[torch.Tensor] = None
        self.conf = torch.zeros(64, dtype=torch.int32, device=self.device)
        self.mout = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.step_index = 0
        self.spin_limit = spin_limit_cycles
        self.kernel_launches = 0
        self._imported = False
        torch.cuda.synchronize(self.device)

    # -- transport ------------------------------------------------------------
    def connect(self) -> None:
        import torch.distributed as dist
        if self.world == 1:
            return
        mine = (native.ipc_export(self.shard), native.ipc_export(self.ctrl))
        handles = [None] * self.world
        dist.all_gather_object(handles, mine)
        for q in range(self.world):
            if q != self.rank:
                self.shards[q] = native.ipc_import(handles[q][0])
                self.ctrls[q] = native.ipc_import(handles[q][1])
        self._imported = True
        dist.barrier()

    def set_peers(self, shards: List[int], ctrls: List[int]) -> None:
        self.shards, self.ctrls = list(shards), list(ctrls)

    def _comm(self) -> native.PsbSparseComm:
        c = native.PsbSparseComm()
        for q in range(self.world):
            c.shard[q] = self.shards[q]
            c.phase[q] = self.ctrls[q]
        c.ticket, c.status = self.ctrl + 128, self.ctrl + 132
        c.rank, c.num_ranks, c.rows_per_shard, c.spin_limit = self.rank, self.world, self.Fs, self.spin_limit
        return c

    # -- the step ---------------------------------------------------------------
    def _launch(self, mode: int, wait: int, publish: int, batch: DeviceCSR) -> None:
        if self.residuals is None or self.residuals.shape[0] < batch.n:
            self.residuals = torch.zeros((batch.n, 8), dtype=torch.float32, device=self.device)
        scale = -self.eta / (batch.n * self.num_workers)
        comm = self._comm()
        native.check(native.lib().psb_sparse_step(
            native.C.byref(comm), mode, wait, publish, batch.indptr.data_ptr(), batch.indices.data_ptr(),
            batch.values.data_ptr(), batch.labels.data_ptr(), batch.n, self.F, self.R, scale,
            self.residuals.data_ptr(), self.loss.data_ptr(), native._stream()), "psb_sparse_step")
        self.kernel_launches += 1

    def step_forward(self, batch: DeviceCSR) -> None:
        self.loss.zero_()
        k = self.step_index
        self._launch(1, 2 * k if (self.world > 1 and k > 0) else 0, 2 * k + 1 if self.world > 1 else 0, batch)

    def step_scatter(self, batch: DeviceCSR) -> None:
        k = self.step_index
        self._launch(2, 2 * k + 1 if self.world > 1 else 0, 2 * k + 2 if self.world > 1 else 0, batch)
        self.step_index += 1

    def step(self, batch: DeviceCSR) -> None:
        """One worker SGD step.  Sequential: forward | barrier | scatter | barrier.  Eventual: one kernel."""
        if self.model == 0:
            self.step_forward(batch)
            self.step_scatter(batch)
        else:
            self.loss.zero_()
            if self.host_delay_s:
                import time
                time.sleep(self.host_delay_s)
            v = self.step_index
            if self.model > 0 and self.world > 1:
                # bounded delay: start step v only when every rank has completed step v - k; publish v + 1 at the end
                self._launch(3, max(0, v - self.model), v + 1, batch)
            else:
                self._launch(3, 0, 0, batch)
            self.step_index += 1

    def completed_steps(self) -> np.ndarray:
        """Steps completed by every rank as this rank sees them (the bounded-delay clock words; zeros otherwise)."""
        return _view(self.ctrl, self.world, self.device, torch.int32).cpu().numpy()

    def mean_loss(self, batch: DeviceCSR) -> float:
        return float(self.loss.item()) / batch.n

    def evaluate(self, test: DeviceCSR):
        comm = self._comm()
        native.check(native.lib().psb_sparse_eval(native.C.byref(comm), test.indptr.data_ptr(), test.indices.data_ptr(),
                                                  test.values.data_ptr(), test.labels.data_ptr(), test.n, self.F, self.R,
                                                  self.conf.data_ptr(), native._stream()), "psb_sparse_eval")
        native.check(native.lib().psb_eval_finalize(self.conf.data_ptr(), self.mout.data_ptr(), native._stream()),
                     "psb_eval_finalize")
        f1, acc = self.mout.cpu().tolist()
        return f1, acc

    def shard_weights(self) -> torch.Tensor:
        return _view(self.shard, (self.Fs + 1) * 8, self.device).view(self.Fs + 1, 8)

    def status(self) -> int:
        return int(_view(self.ctrl + 132, 1, self.device, torch.int32).item())

    def close(self) -> None:
        if self.shard:
            torch.cuda.synchronize(self.device)
            if self._imported:
                for q in range(self.world):
                    if q != self.rank:
                        native.lib().psb_ipc_close(self.shards[q])
                        native.lib().psb_ipc_close(self.ctrls[q])
            native.device_free(self.shard)
            native.device_free(self.ctrl)
            self.shard = 0


class VirtualSparseCluster:
    """N shards on ONE device (tests): forwards of all ranks, then scatters of all ranks, in launch order."""

    def __init__(self, num_features: int, num_ranks: int, consistency_model: int = 0, eta: float = 0.5):
        self.engines = [SparseShardedEngine(num_features, 5, r, num_ranks, num_ranks, eta, consistency_model)
                        for r in range(num_ranks)]
        shards = [e.shard for e in self.engines]
        ctrls = [e.ctrl for e in self.engines]
        for e in self.engines:
            e.set_peers(shards, ctrls)

    def step(self, batches) -> None:
        if self.engines[0].model == 0:
            for e, b in zip(self.engines, batches):
                e.step_forward(b)
            for e, b in zip(self.engines, batches):
                e.step_scatter(b)
        else:
            for e, b in zip(self.engines, batches):
                e.step(b)

    def weights(self) -> np.ndarray:
        F = self.engines[0].F
        rows = [e.shard_weights()[: e.Fs].cpu().numpy() for e in self.engines]
        return np.concatenate(rows)[:F], self.engines[0].shard_weights()[self.engines[0].Fs].cpu().numpy()

    def close(self) -> None:
        for e in self.engines:
            e.close()


class SlabSparseEngine(SparseShardedEngine):
    """The sharded sparse step with contiguous NVLink transfers (``psb_sparse_step_slab``): request lists to the
    owners, reply slabs back (pull), update slabs out (push), owner-side apply.  Sequentially consistent by
    construction (all pulls of a step precede every apply).  ``max_nnz`` bounds the non-zeros of one batch."""

    STAGES = (1, 2, 4, 8)

    def __init__(self, num_features: int, num_classes: int = 5, rank: int = 0, world: int = 1,
                 num_workers_total: Optional[int] = None, eta: float = 0.5, max_nnz: int = 1 << 21,
                 device: Optional[torch.device] = None):
        super().__init__(num_features, num_classes, rank, world, num_workers_total, eta, 0, device)
        self.cap = int(max_nnz)
        N = world
        self._raw = {"req": native.device_malloc(2 * N * self.cap * 4), "req_cnt": native.device_malloc(1024),
                     "rep": native.device_malloc(N * self.cap * 32), "upd": native.device_malloc(N * self.cap * 32)}
        _view(self._raw["req_cnt"], 256, self.device, torch.int32).zero_()
        self.cnt = torch.zeros(16, dtype=torch.int32, device=self.device)
        self.pos = torch.zeros(self.cap, dtype=torch.int32, device=self.device)
        self.peer = {k: [0] * N for k in self._raw}
        for k, v in self._raw.items():
            self.peer[k][rank] = v
        torch.cuda.synchronize(self.device)

    def connect(self) -> None:
        import torch.distributed as dist
        super().connect()
        if self.world == 1:
            return
        mine = {k: native.ipc_export(v) for k, v in self._raw.items()}
        handles = [None] * self.world
        dist.all_gather_object(handles, mine)
        for q in range(self.world):
            if q != self.rank:
                for k in self._raw:
                    self.peer[k][q] = native.ipc_import(handles[q][k])
        dist.barrier()

    def set_slab_peers(self, engines) -> None:
        for k in self._raw:
            self.peer[k] = [e._raw[k] for e in engines]

    def _slab_comm(self) -> native.PsbSlabComm:
        c = native.PsbSlabComm()
        c.base = self._comm()
        for q in range(self.world):
            c.req[q], c.req_cnt[q] = self.peer["req"][q], self.peer["req_cnt"][q]
            c.rep[q], c.upd[q] = self.peer["rep"][q], self.peer["upd"][q]
        c.cnt, c.pos, c.cap = self.cnt.data_ptr(), self.pos.data_ptr(), self.cap
        return c

    def launch_stage(self, batch: DeviceCSR, stage_mask: int) -> None:
        if batch.nnz > self.cap:
            raise ValueError(f"batch has {batch.nnz} non-zeros, the slabs hold {self.cap}")
        if stage_mask & 1:
            self.loss.zero_()
        scale = -self.eta / (batch.n * self.num_workers)
        comm = self._slab_comm()
        native.check(native.lib().psb_sparse_step_slab(
            native.C.byref(comm), self.step_index, batch.indptr.data_ptr(), batch.indices.data_ptr(),
            batch.values.data_ptr(), batch.labels.data_ptr(), batch.n, batch.nnz, self.F, self.R, scale,
            self.loss.data_ptr(), stage_mask, native._stream()), "psb_sparse_step_slab")
        self.kernel_launches += bin(stage_mask).count("1")
        if stage_mask & 8:
            self.step_index += 1

    def step(self, batch: DeviceCSR) -> None:
        self.launch_stage(batch, 15)

    def close(self) -> None:
        if self.shard:
            torch.cuda.synchronize(self.device)
            if self._imported:
                for q in range(self.world):
                    if q != self.rank:
                        for k in self._raw:
                            native.lib().psb_ipc_close(self.peer[k][q])
            for v in self._raw.values():
                native.device_free(v)
        super().close()


class VirtualSlabCluster:
    """N slab engines on ONE device: every stage is launched for all ranks before the next stage."""

    def __init__(self, num_features: int, num_ranks: int, eta: float = 0.5, max_nnz: int = 1 << 16):
        self.engines = [SlabSparseEngine(num_features, 5, r, num_ranks, num_ranks, eta, max_nnz) for r in range(num_ranks)]
        shards, ctrls = [e.shard for e in self.engines], [e.ctrl for e in self.engines]
        for e in self.engines:
            e.set_peers(shards, ctrls)
            e.set_slab_peers(self.engines)

    def step(self, batches) -> None:
        for stage in SlabSparseEngine.STAGES:
            for e, b in zip(self.engines, batches):
                e.launch_stage(b, stage)

    weights = VirtualSparseCluster.weights

    def close(self) -> None:
        for e in self.engines:
            e.close()
