"""Worker engine on SPARSE-row buffers: the reference's local solve (2 L-BFGS iterations, standardisation,
centring — ``LogisticRegressionTaskSpark.java:146-184`` fits on ``Vectors.sparse`` of any width) for models the dense
tcgen05 kernel does not take: ``num_features`` beyond 1024 or not a multiple of 128, or rows so sparse that dense fp16
rings would stream mostly zeros.

Same shape as ``DeviceEngine`` — one object per rank, the workers' input rings, a peer-memory block, ``step()`` = one
fused kernel (``psb_worker_step_csr``: pull, local solve, push, apply) — but a ring slot holds at most ``max_nnz``
(index, fp16 value) pairs (ELL) instead of a dense row.  All consistency models of ``csrc/comm.cuh`` work unchanged,
including per-worker clocks (``async_granularity="worker"``).
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from ..config import PSConfig
from ..data.csr import CSRBatch, pack_stage_rows
from ..models.logreg import LogRegLayout
from ..ops import native
from .device_engine import _view


class SparseRowEngine:
    def __init__(self, cfg: PSConfig, workers_per_rank: int, max_nnz: int, rank: int = 0, world: int = 1,
                 device: Optional[torch.device] = None, w0: Optional[np.ndarray] = None,
                 async_granularity: str = "rank", max_clocks_per_launch: int = 64,
                 spin_limit_cycles: int = 20_000_000_000):
        if not torch.cuda.is_available():
            raise native.NativeError("SparseRowEngine needs a CUDA device")
        native.lib()
        self.cfg, self.W, self.K, self.rank, self.world = cfg, workers_per_rank, int(max_nnz), rank, world
        self.device = dev = device or torch.device("cuda", torch.cuda.current_device())
        self.layout = LogRegLayout(cfg.num_features, cfg.num_classes)
        F, cap = cfg.num_features, cfg.max_buffer_size
        self.F, self.cap, self.nv = F, cap, self.layout.device_size
        W, K = self.W, self.K
        self.idx = torch.zeros((W, cap, K), dtype=torch.int32, device=dev)
        self.val = torch.zeros((W, cap, K), dtype=torch.float16, device=dev)
        self.nnz = torch.zeros((W, cap), dtype=torch.int32, device=dev)
        self.labels = torch.zeros((W, cap), dtype=torch.int32, device=dev)
        self.win_host = torch.zeros((W, 2), dtype=torch.int32).pin_memory()
        self.win = torch.zeros((W, 2), dtype=torch.int32, device=dev)
        self.loss = torch.zeros(W, dtype=torch.float32, device=dev)
        self.loss_host = torch.zeros(W, dtype=torch.float32).pin_memory()
        self.info = torch.zeros((W, 8), dtype=torch.int32, device=dev)
        self.delta = torch.zeros((W, self.nv), dtype=torch.float32, device=dev)
        self.scratch = torch.zeros((W, native.scratch_floats(F)), dtype=torch.float32, device=dev)
        self.solver = native.make_solver(F, self.layout.R, cap, cfg.local_solver, cfg.max_iter, cfg.standardize,
                                         cfg.center, cfg.sgd_lr)
        self.total_workers = W * world
        self.lr = 1.0 / self.total_workers                      # ServerProcessor.java:36
        self.clock, self.spin_limit = 0, spin_limit_cycles
        self.async_granularity = async_granularity
        self.max_clocks_per_launch = max(1, int(max_clocks_per_launch))
        w0 = self.layout.zeros() if w0 is None else np.asarray(w0, dtype=np.float32)
        self.w0 = torch.from_numpy(w0.reshape(-1).copy()).to(dev)
        self.block = native.device_malloc(native.comm_block_bytes(F, world))
        self.peers: List[int] = [0] * world
        self.peers[rank] = self.block
        native.check(native.lib().psb_comm_init_block(self.block, self.w0.data_ptr(), F, world, rank, native._stream()),
                     "psb_comm_init_block")
        self.buffers = [native.NativeBuffer(cfg.min_buffer_size, cap, cfg.buffer_size_coefficient) for _ in range(W)]
        self.hist = torch.zeros((self.max_clocks_per_launch, W, 4), dtype=torch.int32, device=dev)
        self.hist_n = 0
        self.rows_ingested = self.kernel_launches = 0
        self._imported = False
        torch.cuda.synchronize(dev)

    def connect(self) -> None:
        import torch.distributed as dist
        if self.world == 1:
            return
        handles = [None] * self.world
        dist.all_gather_object(handles, native.ipc_export(self.block))
        for q in range(self.world):
            if q != self.rank:
                self.peers[q] = native.ipc_import(handles[q])
        self._imported = True
        dist.barrier()

    @property
    def comm_mode(self) -> int:
        if self.cfg.consistency_model == 0:
            return native.COMM_SEQUENTIAL
        return native.COMM_WORKER if self.async_granularity == "worker" else native.COMM_ASYNC

    # ------------------------------------------------------------------ ingestion
    def ingest(self, batch: CSRBatch, workers: np.ndarray, now_ms: Optional[np.ndarray] = None) -> None:
        """Stream the rows of `batch` into their workers' rings (dynamic-buffer policy on the host, one H2D of the
        packed records, one store kernel).  Rows with more than ``max_nnz`` non-zeros keep their first ``max_nnz``."""
        n = len(batch)
        workers = np.ascontiguousarray(workers[:n], dtype=np.int32)
        now_ms = np.zeros(n, dtype=np.float64) if now_ms is None else np.ascontiguousarray(now_ms[:n], dtype=np.float64)
        ws = np.zeros((n, 2), dtype=np.int32)
        native.buffers_on_batch(self.buffers, workers, now_ms, ws, self.win_host.numpy())
        stage = torch.from_numpy(pack_stage_rows(batch, ws).view(np.int32)).pin_memory()
        stage_dev = stage.to(self.device, non_blocking=True)
        self.win.copy_(self.win_host, non_blocking=True)
        native.check(native.lib().psb_ring_store_ell(stage_dev.data_ptr(), n, stage.shape[1], self.idx.data_ptr(),
                                                     self.val.data_ptr(), self.nnz.data_ptr(), self.labels.data_ptr(),
                                                     self.cap, self.K, self.F, native._stream()), "psb_ring_store_ell")
        torch.cuda.current_stream().synchronize()               # the pinned records may be dropped now
        self.rows_ingested += n
        self.kernel_launches += 1

    # ------------------------------------------------------------------ the step
    def step(self, mode: Optional[int] = None, num_clocks: int = 1) -> None:
        mode = self.comm_mode if mode is None else mode
        c = native.PsbComm()
        c.mode, c.rank, c.num_ranks, c.workers_per_rank = mode, self.rank, self.world, self.W
        c.clock, c.lr, c.spin_limit = self.clock, self.lr, self.spin_limit
        cm = self.cfg.consistency_model
        c.max_delay, c.delay_worker = (cm if cm > 0 else -1), -1
        for q in range(self.world):
            c.peers[q] = self.peers[q]
        if mode == native.COMM_WORKER:
            if not 1 <= num_clocks <= self.max_clocks_per_launch:
                raise ValueError(f"num_clocks must be in [1, {self.max_clocks_per_launch}]")
            c.num_clocks, c.worker_base, c.total_workers = num_clocks, self.rank * self.W, self.total_workers
            c.hist = self.hist.data_ptr()
            self.hist_n = num_clocks
        elif num_clocks != 1:
            raise ValueError("several clocks per launch need async_granularity='worker' and consistency_model != 0")
        none = mode == native.COMM_NONE
        native.check(native.lib().psb_worker_step_csr(
            native.C.byref(self.solver), native.C.byref(c), self.W, self.K, self.idx.data_ptr(), self.val.data_ptr(),
            self.nnz.data_ptr(), self.labels.data_ptr(), self.win.data_ptr(), self.w0.data_ptr() if none else None,
            self.delta.data_ptr() if none else None, self.loss.data_ptr(), self.info.data_ptr(), self.scratch.data_ptr(),
            native._stream()), "psb_worker_step_csr")
        self.kernel_launches += 1
        self.clock += num_clocks

    def read_loss(self) -> torch.Tensor:
        self.loss_host.copy_(self.loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.loss_host

    def weights(self) -> torch.Tensor:
        L = native.lib()
        if self.comm_mode == native.COMM_SEQUENTIAL:
            ptr = L.psb_comm_replica(self.block, self.F, self.world, self.clock & 1)
            return _view(ptr, self.nv, self.device).view(self.layout.device_rows, 8)
        out = torch.empty(self.nv, dtype=torch.float32, device=self.device)
        slice_f = ((self.nv // 4 + self.world - 1) // self.world) * 4
        for q in range(self.world):
            b, e = min(q * slice_f, self.nv), min((q + 1) * slice_f, self.nv)
            if e > b:
                out[b:e].copy_(_view(L.psb_comm_shard_weights(self.peers[q], self.F, self.world), e - b, self.device))
        return out.view(self.layout.device_rows, 8)

    def status(self) -> int:
        return int(_view(native.lib().psb_comm_status(self.block, self.F, self.world), 1, self.device,
                         torch.int32).item())

    def close(self) -> None:
        if self.block:
            torch.cuda.synchronize(self.device)
            if self._imported:
                for q, p in enumerate(self.peers):
                    if q != self.rank and p:
                        native.lib().psb_ipc_close(p)
            native.device_free(self.block)
            self.block = 0
