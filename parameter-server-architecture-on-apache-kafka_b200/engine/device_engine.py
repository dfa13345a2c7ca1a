"""Device engine: the fused parameter-server step on one GPU rank.

One object per rank owns what the reference spreads over a worker JVM, a server
JVM and three Kafka topics:

* the workers' input rings (``INPUT_DATA_BUFFER`` state store, ``WorkerApp.java:40-42``)
  as one dense fp16 tensor ``X[W, cap, F]`` + labels + (head, count) windows;
* this rank's peer-memory block (weights replicas, shard, inboxes, clocks —
  ``csrc/common.cuh: CommLayout``) which stands in for ``WEIGHTS_TOPIC`` /
  ``GRADIENTS_TOPIC`` / ``MessageTracker``;
* a pinned staging ring + side stream for ingestion (``CsvProducer`` -> ``INPUT_DATA``).

``step()`` enqueues ONE kernel: pull (gated by the consistency model) -> local
solve -> push -> apply (``psb_worker_step``).  Nothing is synchronised with the
host unless the caller asks for the loss.

Asynchrony comes in two granularities (``async_granularity``).  ``"rank"``: the
workers of a GPU advance in lockstep and the relaxed models hold between GPUs.
``"worker"``: every worker CTA owns a clock word and is gated on its own, as the
reference's ``MessageTracker`` does per partition (``MessageTracker.java:42-88``);
``step(num_clocks=n)`` then lets each worker run n clocks at its own pace inside
one launch, and ``groups > 1`` launches the workers of a GPU as several
independent groups, each on its own stream (``step_group``), so that a slow
group never holds back a launch boundary of a fast one.

Multi-GPU: one engine per process, blocks exchanged with ``cudaIpc`` through
``torch.distributed`` (``connect()``).  Tests on fewer GPUs than ranks build several
engines on one device (``VirtualCluster``) and step them in clock order, which
never makes one launch wait on another.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from ..config import PSConfig
from ..data.csr import CSRBatch
from ..models.logreg import LogRegLayout
from ..ops import native


class DeviceEngine:
    def __init__(self, cfg: PSConfig, workers_per_rank: int, rank: int = 0, world: int = 1, impl: int = native.IMPL_SIMT,
                 device: Optional[torch.device] = None, w0: Optional[np.ndarray] = None, stage_rows: int = 0,
                 stage_sets: int = 3, spin_limit_cycles: int = 20_000_000_000, async_granularity: str = "rank",
                 groups: int = 1, max_clocks_per_launch: int = 64, external_block: Optional[int] = None,
                 home_remote: bool = False, ring_spare_rows: int = 0, stats_at_ingest: Optional[bool] = None):
        if not torch.cuda.is_available():
            raise native.NativeError("DeviceEngine needs a CUDA device")
        if async_granularity not in ("rank", "worker"):
            raise ValueError("async_granularity must be 'rank' or 'worker'")
        if groups < 1 or workers_per_rank % groups:
            raise ValueError(f"groups ({groups}) must divide workers_per_rank ({workers_per_rank})")
        if groups > 1 and cfg.consistency_model > 0:
            # gated workers wait on each other inside the kernel: they must be ONE launch per GPU (nothing guarantees
            # that separate launches run at the same time)
            raise ValueError("bounded delay runs the workers of a GPU as one launch: groups must be 1")
        native.lib()
        self.cfg, self.W, self.rank, self.world, self.impl = cfg, workers_per_rank, rank, world, impl
        self.async_granularity, self.groups, self.home_remote = async_granularity, groups, bool(home_remote)
        self.max_clocks_per_launch = max(1, int(max_clocks_per_launch))
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.layout = LogRegLayout(cfg.num_features, cfg.num_classes)
        # Ring of every worker: the largest window (max_buffer_size) plus `ring_spare_rows` spare slots.  New rows are
        # written at the head, i.e. into slots OUTSIDE the current window as long as a batch brings at most
        # `ring_spare_rows` rows per worker: such a batch may be stored while a step still trains on the current window
        # (ingest_commit(overlap=True)).  Kept a multiple of 128 rows so that the tcgen05 kernel's slabs stay whole.
        spare = (int(ring_spare_rows) + 127) // 128 * 128 if ring_spare_rows > 0 else 0
        F, cap = cfg.num_features, cfg.max_buffer_size + spare
        self.F, self.cap, self.nv, self.ring_spare = F, cap, self.layout.device_size, spare
        dev = self.device
        self.X = torch.zeros((self.W, cap, F), dtype=torch.float16, device=dev)
        self.labels = torch.zeros((self.W, cap), dtype=torch.int32, device=dev)
        self.win_host = torch.zeros((self.W, 2), dtype=torch.int32).pin_memory()
        # the windows live twice on the device: a step reads the table that the last commit wrote, an overlapped
        # commit writes the other one
        self._win = [torch.zeros((self.W, 2), dtype=torch.int32, device=dev) for _ in range(2)]
        self._win_i = 0
        # Standardisation statistics kept current at ingestion time (psb_window_stats_update) when the rings have spare
        # slots (the rows that leave a window are then still in the ring when the batch that pushes them out has been
        # stored); otherwise the worker kernel computes them itself, fused into its first pass.  Tables flip with `win`.
        # (pays when clocks outnumber arriving batches — the reference's regime: seconds between rows, the buffers
        # unchanged for most clocks; with a batch per clock the fused pass is as cheap.)  Default: on with spare slots.
        want = spare > 0 if stats_at_ingest is None else bool(stats_at_ingest)
        self.stats_at_ingest = bool(want and spare > 0 and cfg.standardize and F % 4 == 0)
        if self.stats_at_ingest:
            self.stat_sums = torch.zeros((self.W, F, 2), dtype=torch.float64, device=dev)
            self._inv = [torch.zeros((self.W, F), dtype=torch.float32, device=dev) for _ in range(2)]
        self._commits_since_step = 0
        self._step_done = [torch.cuda.Event(), torch.cuda.Event()]    # completion of the last two steps
        self._steps = 0
        self.loss = torch.zeros(self.W, dtype=torch.float32, device=dev)
        self.info = torch.zeros((self.W, 8), dtype=torch.int32, device=dev)
        self.delta = torch.zeros((self.W, self.nv), dtype=torch.float32, device=dev)
        self.scratch = torch.zeros((self.W, native.scratch_floats(F)), dtype=torch.float32, device=dev)
        self.loss_host = torch.zeros(self.W, dtype=torch.float32).pin_memory()
        self.solver = native.make_solver(F, self.layout.R, cap, cfg.local_solver, cfg.max_iter, cfg.standardize,
                                         cfg.center, cfg.sgd_lr)
        self.total_workers = self.W * world
        self.lr = 1.0 / self.total_workers                      # ServerProcessor.java:36
        self.clock = 0
        self.spin_limit = spin_limit_cycles
        w0 = self.layout.zeros() if w0 is None else np.asarray(w0, dtype=np.float32)
        self.w0 = torch.from_numpy(w0.reshape(-1).copy()).to(dev)
        if self.total_workers > native.MAX_WORKERS and async_granularity == "worker":
            raise ValueError(f"per-worker clocks support at most {native.MAX_WORKERS} workers")
        # peer-memory block (raw cudaMalloc so it can be exported over IPC); `external_block`: the block was
        # allocated and initialised by another process (the server app) and is only mapped here
        self.block_bytes = native.comm_block_bytes(F, world)
        self.owns_block = external_block is None
        self.block = native.device_malloc(self.block_bytes) if self.owns_block else int(external_block)
        self.peers: List[int] = [0] * world
        self.peers[rank] = self.block
        if self.owns_block:
            native.check(native.lib().psb_comm_init_block(self.block, self.w0.data_ptr(), F, world, rank,
                                                          native._stream()), "psb_comm_init_block")
        # per-worker asynchrony: one clock per launch group, a record of every (clock, worker) of a launch
        self.gsize = self.W // groups
        self.group_clock = [0] * groups
        self.group_streams = [torch.cuda.Stream(device=dev) for _ in range(groups)] if groups > 1 else [None]
        self.hist = None
        self.delay_worker, self.delay_cycles = -1, 0
        self.publish = False                  # sequential, one rank: keep the block's published snapshot current
        # ingestion: host-side window bookkeeping + a ring of staging sets.  begin() fills set `head` over the
        # copy stream, commit() drains set `tail` into the rings over the ingest stream; with S sets the H2D of
        # batch k+S-1 overlaps the step that trains on batch k.
        self.buffers = [native.NativeBuffer(cfg.min_buffer_size, cfg.max_buffer_size, cfg.buffer_size_coefficient,
                                            ring=cap) for _ in range(self.W)]
        self._buf_handles = None
        self.rows_ingested = 0
        self.stage_rows = stage_rows or self.W * 128
        S = self.stage_sets = max(2, int(stage_sets))
        self.stage_ws_host = [torch.zeros((self.stage_rows, 2), dtype=torch.int32).pin_memory() for _ in range(S)]
        self.stage_win_host = [torch.zeros((self.W, 2), dtype=torch.int32).pin_memory() for _ in range(S)]
        self.stage_change_host = [torch.zeros((self.W, 4), dtype=torch.int32).pin_memory() for _ in range(S)]
        self.stage_change_dev = [torch.zeros((self.W, 4), dtype=torch.int32, device=dev) for _ in range(S)]
        self.stage_dev = [torch.zeros((self.stage_rows, F), dtype=torch.float16, device=dev) for _ in range(S)]
        self.stage_lab_dev = [torch.zeros(self.stage_rows, dtype=torch.int32, device=dev) for _ in range(S)]
        self.stage_ws_dev = [torch.zeros((self.stage_rows, 2), dtype=torch.int32, device=dev) for _ in range(S)]
        self.stage_bitmap_dev = self.stage_off_dev = None            # allocated on the first packed batch
        self.stage_csr_dev = [None] * S                              # (device rows, stride) of an ingest_sparse batch
        self.stage_future = [None] * S                               # slot assignment of the set (helper thread)
        self._slot_thread = None
        self.stage_kind = ["dense"] * S
        self.stage_n = [0] * S
        self.stage_max_rows = [0] * S
        self.stage_head = self.stage_tail = self.stage_pending = 0
        self.h2d_done = [torch.cuda.Event() for _ in range(S)]       # copies of set j have landed
        self.store_done = [torch.cuda.Event() for _ in range(S)]     # set j has been drained into the rings
        self.copy_stream = torch.cuda.Stream(device=dev)
        self.ingest_stream = torch.cuda.Stream(device=dev)
        self.ingest_done = torch.cuda.Event()
        self.h2d_bytes = 0
        self.kernel_launches = 0
        # evaluation buffers
        self.conf = torch.zeros(64, dtype=torch.int32, device=dev)
        self.metrics_dev = torch.zeros(2, dtype=torch.float32, device=dev)
        self.metrics_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        self.test_X = self.test_y = None
        torch.cuda.synchronize(dev)

    # ------------------------------------------------------------------ transport set-up
    def connect(self) -> None:
        """Exchange IPC handles of the peer blocks (one engine per process, torch.distributed up)."""
        import torch.distributed as dist
        if self.world == 1:
            return
        handle = native.ipc_export(self.block)
        handles = [None] * self.world
        dist.all_gather_object(handles, handle)
        for q in range(self.world):
            if q != self.rank:
                self.peers[q] = native.ipc_import(handles[q])
        self._imported = True
        dist.barrier()

    def set_peers(self, peers: List[int]) -> None:
        self.peers = list(peers)

    def _comm(self, mode: int) -> native.PsbComm:
        c = native.PsbComm()
        c.mode, c.rank, c.num_ranks, c.workers_per_rank = mode, self.rank, self.world, self.W
        c.clock, c.lr, c.spin_limit = self.clock, self.lr, self.spin_limit
        cm = self.cfg.consistency_model
        c.max_delay = cm if cm > 0 else -1
        c.home_remote = int(self.home_remote)
        c.publish = int(self.publish and mode == native.COMM_SEQUENTIAL and self.world == 1)
        c.delay_worker = -1
        for q in range(self.world):
            c.peers[q] = self.peers[q]
        return c

    @property
    def comm_mode(self) -> int:
        if self.cfg.consistency_model == 0:
            return native.COMM_SEQUENTIAL
        return native.COMM_WORKER if self.async_granularity == "worker" else native.COMM_ASYNC

    def inject_delay(self, global_worker: int, cycles: int) -> None:
        """Straggler injection for the per-worker modes: that worker sleeps `cycles` SM cycles before every clock."""
        self.delay_worker, self.delay_cycles = int(global_worker), int(cycles)

    def max_resident_workers(self) -> int:
        """Worker CTAs of this shape that fit on the device at once (gated per-worker launches need all W resident)."""
        return int(native.check(native.lib().psb_worker_max_resident(self.solver, self.impl, self.W),
                                "psb_worker_max_resident"))

    # ------------------------------------------------------------------ ingestion
    def load_test(self, test: CSRBatch) -> None:
        self.test_X = torch.from_numpy(test.to_dense(np.float16)).to(self.device)
        self.test_y = torch.from_numpy(test.labels.astype(np.int32)).to(self.device)

    @property
    def win(self) -> torch.Tensor:
        """Device table of the workers' (head, count) windows that the next step will read."""
        return self._win[self._win_i]

    def fill(self, rows: torch.Tensor, labels: torch.Tensor) -> None:
        """Bulk-load every ring: rows fp16 [W, n, F] (device or host), labels int32 [W, n]; n <= max_buffer_size."""
        n = rows.shape[1]
        if n > self.cfg.max_buffer_size:
            raise ValueError(f"fill: {n} rows per worker exceed max_buffer_size {self.cfg.max_buffer_size}")
        self._wait_slot_thread()
        self.X[:, :n].copy_(rows)
        self.labels[:, :n].copy_(labels)
        self.buffers = [native.NativeBuffer(self.cfg.min_buffer_size, self.cfg.max_buffer_size,
                                            self.cfg.buffer_size_coefficient, ring=self.cap) for _ in range(self.W)]
        self._buf_handles = None
        for b in self.buffers:                      # replay n unthrottled arrivals so the host mirror agrees
            for _ in range(n):
                b.on_event(0.0)
        self.win_host[:, 0] = n % self.cap
        self.win_host[:, 1] = n
        for t in self._win:
            t.copy_(self.win_host)
        self._recompute_stats()

    def ingest(self, rows_f16, labels: Optional[torch.Tensor], workers: np.ndarray,
               now_ms: Optional[np.ndarray] = None, wait: bool = True) -> None:
        """Stream `n` new rows (pinned fp16 [n, F] + int32 labels, or a ``native.PackedRows``) to their workers' rings.

        Host: slot per row from the dynamic-buffer policy; device: one H2D of the rows, one of the
        (worker, slot) table, one store kernel — on the copy / ingestion streams.
        """
        self.ingest_begin(rows_f16, labels, workers, now_ms)
        self.ingest_commit(wait)

    def ingest_sparse(self, batch: CSRBatch, workers: np.ndarray, now_ms: Optional[np.ndarray] = None) -> None:
        """ingest() for CSR rows: ships (index, value) records and lets ``psb_ring_scatter`` expand them on the
        device (fewer PCIe bytes than dense fp16 below ~25 % density; see PackedRows for the denser case)."""
        from ..data.csr import pack_stage_rows
        n = len(batch)
        j = self._begin_set(n, workers, now_ms, sync=True)
        stage = torch.from_numpy(pack_stage_rows(batch, self.stage_ws_host[j].numpy()[:n]).view(np.int32)).pin_memory()
        with torch.cuda.stream(self.copy_stream):
            stage_dev = stage.to(self.device, non_blocking=True)
            self.h2d_done[j].record(self.copy_stream)
        stage_dev.record_stream(self.ingest_stream)
        self.stage_csr_dev[j] = (stage_dev, stage.shape[1], stage)     # pinned source must outlive the async copy
        self._end_set(j, n, "csr", stage.numel() * 4 + self.W * 8)
        self.ingest_commit()

    def _begin_set(self, n: int, workers: np.ndarray, now_ms: Optional[np.ndarray], sync: bool = False) -> int:
        """Claim staging set `head` and hand its slot assignment (dynamic-buffer policy: ring slot per row, the
        windows they produce) to the engine's helper thread, which fills the set's pinned tables.  The tables are
        needed only at commit time, so the ~0.4 ms of host work per 38k rows overlaps the caller's enqueueing and
        the running step; batches are assigned strictly in begin order (one helper thread)."""
        assert n <= self.stage_rows, "ingest batch larger than a staging set"
        if self.stage_pending == self.stage_sets:
            raise native.NativeError("ingest_begin: every staging set holds an uncommitted batch; call ingest_commit")
        j = self.stage_head
        self.store_done[j].synchronize()                    # the set's pinned tables and device buffers are free again
        if self._buf_handles is None:
            self._buf_handles = native.buffer_handles(self.buffers)
        workers = np.ascontiguousarray(workers[:n], dtype=np.int32)
        now_ms = np.zeros(n, dtype=np.float64) if now_ms is None else np.ascontiguousarray(now_ms[:n], dtype=np.float64)
        handles, buffers = self._buf_handles, self.buffers

        def assign():
            # slot of every row, the windows after the batch, what changed per worker (for the statistics update) and
            # the most rows any worker receives (decides whether the store may overlap a running step): one native call
            self.stage_max_rows[j] = native.buffers_on_batch(
                buffers, workers, now_ms, self.stage_ws_host[j].numpy(), self.win_host.numpy(), handles=handles,
                change=self.stage_change_host[j].numpy())
            self.stage_win_host[j].copy_(self.win_host)

        if self._slot_thread is None:
            from concurrent.futures import ThreadPoolExecutor
            self._slot_thread = ThreadPoolExecutor(max_workers=1, thread_name_prefix="psb-slots")
        self.stage_future[j] = self._slot_thread.submit(assign)
        if sync:
            self.stage_future[j].result()
        return j

    def _wait_slot_thread(self) -> None:
        """Block until the helper thread has finished every slot assignment handed to it."""
        for f in self.stage_future:
            if f is not None:
                f.result()

    def _end_set(self, j: int, n: int, kind: str, nbytes: int) -> None:
        self.stage_n[j], self.stage_kind[j] = n, kind
        self.stage_head = (j + 1) % self.stage_sets
        self.stage_pending += 1
        self.rows_ingested += n
        self.h2d_bytes += nbytes

    def ingest_begin(self, rows_f16, labels: Optional[torch.Tensor], workers: np.ndarray,
                     now_ms: Optional[np.ndarray] = None) -> None:
        """First half of ingest(): slot assignment + host->device copies into a staging set.  Does not touch
        the rings, so it may run while a step is training on them; up to ``stage_sets`` batches may be begun
        ahead of their ingest_commit (pipelined ingestion).

        ``rows_f16`` is either pinned fp16 [n, F] (with ``labels``) or a ``native.PackedRows`` batch (bitmap +
        fp16 values; ``labels`` ignored) — the packed form moves 128 B + 2 B per non-zero per row over PCIe."""
        packed = isinstance(rows_f16, native.PackedRows)
        n = rows_f16.n if packed else rows_f16.shape[0]
        j = self._begin_set(n, workers, now_ms)
        s = self.copy_stream
        if packed:
            pk = rows_f16
            assert pk.F == self.F
            if self.stage_bitmap_dev is None:
                self.stage_bitmap_dev = [torch.zeros((self.stage_rows, self.F // 32), dtype=torch.int32,
                                                     device=self.device) for _ in range(self.stage_sets)]
                self.stage_off_dev = [torch.zeros(self.stage_rows + 1, dtype=torch.int32, device=self.device)
                                      for _ in range(self.stage_sets)]
            if pk.palette:
                assert pk.payload_bytes <= self.stage_rows * self.F * 2, "palette payload larger than a staging set"
                body = (self.stage_dev[j], pk.payload, pk.payload_bytes)
            else:
                body = (self.stage_dev[j], pk.values, pk.nnz * 2)
            copies = [body, (self.stage_bitmap_dev[j], pk.bitmap, n * (self.F // 32) * 4),
                      (self.stage_off_dev[j], pk.row_off, (n + 1) * 4), (self.stage_lab_dev[j], pk.labels, n * 4)]
            nbytes = pk.nbytes
        else:
            if rows_f16.dtype != torch.float16 or labels.dtype != torch.int32 or not rows_f16.is_contiguous():
                raise native.NativeError("ingest: rows must be contiguous fp16 [n, F] and labels int32 (pinned host tensors)")
            copies = [(self.stage_dev[j], rows_f16, n * self.F * 2), (self.stage_lab_dev[j], labels, n * 4)]
            nbytes = n * self.F * 2 + n * 4
        native.h2d_copies(copies, stream=s)                 # one native call: rows / payload, bitmap, offsets, labels
        self.h2d_done[j].record(s)
        self._end_set(j, n, ("palette" if rows_f16.palette else "bitmap") if packed else "dense",
                      nbytes + n * 8 + self.W * 8)

    def ingest_from_device(self, n: int, workers: np.ndarray, now_ms: Optional[np.ndarray], bitmap: int, row_off: int,
                           payload: int, labels: int, wait: bool = True):
        """ingest() for a palette-packed batch that ALREADY sits in device memory (raw device addresses; e.g. a slot
        of the INPUT_DATA topic written by the server process, ``psb200.parallel.p2p``): slot assignment on the host,
        one copy of the (worker, slot) table, one store kernel that reads the batch in place.  Returns the ingestion
        stream so that the caller can order its 'slot consumed' signal behind the store."""
        if self.stage_pending:
            raise native.NativeError("ingest_from_device: commit the begun batches first")
        j = self._begin_set(n, workers, now_ms, sync=True)
        s = self.ingest_stream
        s.wait_stream(torch.cuda.current_stream())
        for gs in self.group_streams:
            if gs is not None:
                s.wait_stream(gs)
        self._win_i ^= 1
        self._commits_since_step += 1
        native.h2d_copies([(self.win, self.stage_win_host[j], self.W * 8),
                           (self.stage_ws_dev[j], self.stage_ws_host[j], n * 8)], stream=s)
        native.ring_store_palette(int(bitmap), int(row_off), int(payload), int(labels), self.stage_ws_dev[j], n, self.X,
                                  self.labels, self.cap, self.F, stream=s)
        self._update_stats(j, s)
        self.store_done[j].record(s)
        self.ingest_done.record(s)
        self.kernel_launches += 1
        self.stage_head = (j + 1) % self.stage_sets
        self.stage_tail = self.stage_head
        self.rows_ingested += n
        self.h2d_bytes += n * 8 + self.W * 8
        if wait:
            torch.cuda.current_stream().wait_event(self.ingest_done)
        return s

    def _update_stats(self, j: int, s) -> None:
        """Behind the store of staging set j (stream s): subtract the rows that left each window, add the rows that
        entered it, refresh 1/std in the table that flipped with `win`."""
        if not self.stats_at_ingest:
            return
        native.h2d_copies([(self.stage_change_dev[j], self.stage_change_host[j], self.W * 16)], stream=s)
        native.window_stats_update(self.X, self.stage_change_dev[j], self.W, self.cap, self.F, self.stat_sums,
                                   self._inv[self._win_i], stream=s)
        self.kernel_launches += 1

    def _recompute_stats(self) -> None:
        """Statistics of every window from scratch (after fill / load_state_dict), into both tables."""
        if not self.stats_at_ingest:
            return
        ch = torch.zeros((self.W, 4), dtype=torch.int32)
        ch[:, 0], ch[:, 2], ch[:, 3] = self.win_host[:, 0], -1, self.win_host[:, 1]
        ch = ch.to(self.device)
        for t in self._inv:
            native.window_stats_update(self.X, ch, self.W, self.cap, self.F, self.stat_sums, t)
        torch.cuda.current_stream().synchronize()

    def ingest_commit(self, wait: bool = True, overlap: bool = False) -> None:
        """Second half: store the oldest begun batch into the rings — after the step that is running (default), or,
        with ``overlap=True``, WHILE the most recently enqueued step runs: the new rows go into ring slots outside that
        step's windows (needs ``ring_spare_rows`` >= the rows any worker receives from the batch, one overlapped commit
        per step; otherwise the call quietly orders itself behind the step).  The next step waits for the store."""
        if self.stage_pending == 0:
            return
        j = self.stage_tail
        n = self.stage_n[j]
        s = self.ingest_stream
        torch.cuda.nvtx.range_push("psb.ingest_commit")
        self.stage_future[j].result()                       # slot tables of this batch are written (helper thread)
        overlap = (overlap and self.groups == 1 and self._commits_since_step == 0 and self._steps >= 1
                   and 0 < self.stage_max_rows[j] <= self.ring_spare)
        if overlap:
            # the slots written now were inside the window of the step BEFORE the running one: wait for that one only
            if self._steps >= 2:
                s.wait_event(self._step_done[self._steps & 1])
        else:
            s.wait_stream(torch.cuda.current_stream())      # the previous step must be done with the rings
            for gs in self.group_streams:                   # ... on whichever stream its groups ran
                if gs is not None:
                    s.wait_stream(gs)
        s.wait_event(self.h2d_done[j])
        self._win_i ^= 1                                    # the running step keeps reading the other table
        self._commits_since_step += 1
        tables = [(self.win, self.stage_win_host[j], self.W * 8)]
        if self.stage_kind[j] != "csr":
            tables.append((self.stage_ws_dev[j], self.stage_ws_host[j], n * 8))
        native.h2d_copies(tables, stream=s)
        if self.stage_kind[j] == "palette":
            native.ring_store_palette(self.stage_bitmap_dev[j], self.stage_off_dev[j], self.stage_dev[j],
                                      self.stage_lab_dev[j], self.stage_ws_dev[j], n, self.X, self.labels, self.cap,
                                      self.F, stream=s)
        elif self.stage_kind[j] == "bitmap":
            native.ring_store_bitmap(self.stage_bitmap_dev[j], self.stage_off_dev[j], self.stage_dev[j],
                                     self.stage_lab_dev[j], self.stage_ws_dev[j], n, self.X, self.labels, self.cap,
                                     self.F, stream=s)
        elif self.stage_kind[j] == "csr":
            stage_dev, stride, _ = self.stage_csr_dev[j]
            native.ring_scatter(stage_dev, n, stride, self.X, self.labels, self.cap, self.F, stream=s)
        else:
            native.ring_store_dense(self.stage_dev[j], self.stage_lab_dev[j], self.stage_ws_dev[j], n, self.X,
                                    self.labels, self.cap, self.F, stream=s)
        self._update_stats(j, s)
        self.store_done[j].record(s)
        self.ingest_done.record(s)
        self.kernel_launches += 1
        self.stage_tail = (j + 1) % self.stage_sets
        self.stage_pending -= 1
        if wait:
            torch.cuda.current_stream().wait_event(self.ingest_done)
        torch.cuda.nvtx.range_pop()

    # ------------------------------------------------------------------ the step
    def step(self, mode: Optional[int] = None, num_clocks: int = 1) -> None:
        """Enqueue one parameter-server clock (one kernel); per-worker mode: `num_clocks` clocks of every group.
        NVTX range 'psb.step' for timeline tools."""
        torch.cuda.nvtx.range_push("psb.step")
        try:
            self._step(mode, num_clocks)
        finally:
            torch.cuda.nvtx.range_pop()

    def _step(self, mode: Optional[int] = None, num_clocks: int = 1) -> None:
        mode = self.comm_mode if mode is None else mode
        if mode == native.COMM_WORKER:
            for g in range(self.groups):
                self.step_group(g, num_clocks, stream=False if self.groups == 1 else None)
            self.clock = min(self.group_clock)
            return
        if num_clocks != 1 and mode != native.COMM_SEQUENTIAL:
            raise ValueError("several clocks per launch need the sequential model or async_granularity='worker'")
        comm = self._comm(mode)
        comm.num_clocks = int(num_clocks)
        self.solver.inv_std = self._inv[self._win_i].data_ptr() if self.stats_at_ingest else None
        native.worker_step(self.solver, comm, self.W, self.impl, self.X, self.labels, self.win,
                           self.w0 if mode == native.COMM_NONE else None,
                           self.delta if mode == native.COMM_NONE else None,
                           self.loss, self.info, self.scratch)
        self._step_done[self._steps & 1].record()
        self._steps += 1
        self._commits_since_step = 0
        self.kernel_launches += 1
        self.clock += int(num_clocks)

    def step_group(self, g: int, num_clocks: int = 1, stream=None) -> None:
        """Per-worker mode: enqueue `num_clocks` clocks of the workers of group g — one launch on the group's own
        stream (``stream=False``: the current stream).  Every worker of the group runs its clocks back to back at its
        own pace, gated against ALL workers' clock words; a (loss, clock, lag, lead) record per clock and worker is
        kept for ``read_hist``."""
        if self.comm_mode != native.COMM_WORKER:
            raise native.NativeError("step_group needs async_granularity='worker' and consistency_model != 0")
        n = int(num_clocks)
        if not 1 <= n <= self.max_clocks_per_launch:
            raise ValueError(f"num_clocks must be in [1, {self.max_clocks_per_launch}]")
        if self.hist is None:
            self.hist = torch.zeros((self.groups, self.max_clocks_per_launch, self.gsize, 4), dtype=torch.int32,
                                    device=self.device)
            self.hist_n = [0] * self.groups
            if self.cfg.consistency_model > 0 and self.W > self.max_resident_workers():
                raise native.NativeError(
                    f"bounded delay with per-worker clocks: {self.W} workers per GPU, but only "
                    f"{self.max_resident_workers()} worker CTAs are resident at once (they wait on each other)")
        gs, lo = self.gsize, g * self.gsize
        comm = self._comm(native.COMM_WORKER)
        comm.clock, comm.num_clocks = self.group_clock[g], n
        comm.worker_base, comm.total_workers = self.rank * self.W + lo, self.total_workers
        comm.delay_worker, comm.delay_cycles = self.delay_worker, self.delay_cycles
        comm.hist = self.hist[g].data_ptr()
        self.solver.inv_std = self._inv[self._win_i][lo:lo + gs].data_ptr() if self.stats_at_ingest else None
        st = torch.cuda.current_stream() if stream is False else (stream or self.group_streams[g]
                                                                  or torch.cuda.current_stream())
        if st != torch.cuda.current_stream():
            st.wait_stream(torch.cuda.current_stream())     # fills / ring stores ordered on the caller's stream
        native.worker_step(self.solver, comm, gs, self.impl, self.X[lo:lo + gs], self.labels[lo:lo + gs],
                           self.win[lo:lo + gs], None, None, self.loss[lo:lo + gs], self.info[lo:lo + gs],
                           self.scratch[lo:lo + gs], stream=st)
        self.hist_n[g] = n
        if self.groups == 1:                                # (overlapped ring stores order themselves on these)
            self._step_done[self._steps & 1].record(st)
            self._steps += 1
        self._commits_since_step = 0
        self.kernel_launches += 1
        self.group_clock[g] += n

    def read_hist(self, g: int = 0) -> np.ndarray:
        """Records of group g's last launch as a structured array [num_clocks, group size] with fields loss, clock,
        lag (clock - slowest worker's clock when the gate opened) and lead (fastest worker's clock - clock).
        Synchronises the device."""
        torch.cuda.synchronize(self.device)
        raw = self.hist[g, : self.hist_n[g]].cpu().numpy()
        return raw.view(native.CLOCK_RECORD_DTYPE).reshape(self.hist_n[g], self.gsize)

    def worker_clocks(self) -> np.ndarray:
        """This rank's copy of the per-worker clock table (clocks completed by every worker of every rank)."""
        ptr = native.lib().psb_comm_worker_clocks(self.block, self.F, self.world)
        return _view(ptr, self.total_workers, self.device, torch.int32).cpu().numpy()

    def read_loss(self) -> torch.Tensor:
        """Device -> pinned host copy of the per-worker losses of the last step (synchronises)."""
        self.loss_host.copy_(self.loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return self.loss_host

    def read_loss_begin(self) -> int:
        """Enqueue the device -> pinned host copy of the losses of the step just enqueued; returns a ticket for
        ``read_loss_end``.  Lets the host read step k's result while step k+1 is already queued (two tickets may be
        outstanding)."""
        k = self._loss_ticket = (getattr(self, "_loss_ticket", -1) + 1) % 2
        if not hasattr(self, "_loss_ring"):
            self._loss_ring = [torch.zeros(self.W, dtype=torch.float32).pin_memory() for _ in range(2)]
            self._loss_ev = [torch.cuda.Event() for _ in range(2)]
        self._loss_ring[k].copy_(self.loss, non_blocking=True)
        self._loss_ev[k].record()
        return k

    def read_loss_end(self, ticket: int) -> torch.Tensor:
        """Wait for the copy behind ``ticket`` and return the pinned per-worker losses."""
        self._loss_ev[ticket].synchronize()
        return self._loss_ring[ticket]

    def weights(self) -> torch.Tensor:
        """Current global weights as seen by this rank ([(F+1), 8] view of device memory)."""
        L = native.lib()
        if self.comm_mode == native.COMM_SEQUENTIAL:
            ptr = L.psb_comm_replica(self.block, self.F, self.world, self.clock & 1)
            return _view(ptr, self.nv, self.device).view(self.layout.device_rows, 8)
        if self.world == 1:
            ptr = L.psb_comm_shard_weights(self.block, self.F, self.world)
            return _view(ptr, self.nv, self.device).view(self.layout.device_rows, 8)
        # relaxed models on several ranks: the authoritative weights are the owners' shards — read each over its
        # peer mapping into one local tensor (an unsynchronised snapshot, like a reply of the reference's server
        # in these models: ServerProcessor.java:172-182 sends whatever has been applied so far)
        out = torch.empty(self.nv, dtype=torch.float32, device=self.device)
        slice_f = ((self.nv // 4 + self.world - 1) // self.world) * 4
        for q in range(self.world):
            b, e = min(q * slice_f, self.nv), min((q + 1) * slice_f, self.nv)
            if e > b:
                src = L.psb_comm_shard_weights(self.peers[q], self.F, self.world)
                out[b:e].copy_(_view(src, e - b, self.device))
        return out.view(self.layout.device_rows, 8)

    def worker_local_weights(self, j: int) -> torch.Tensor:
        """Weights of worker j's locally fitted model of the LAST step (pulled weights + its delta) — what the
        reference evaluates for the worker log row (LogisticRegressionTaskSpark.java:186).  Sequential mode."""
        nv = self.nv
        delta = self.scratch[j, 5 * nv: 6 * nv]              # solver scratch vector 5 = delta (csrc/worker_common.cuh)
        w0 = self.scratch[j, 6 * nv: 7 * nv]                 # vector 6 = the weights this step started from
        return (w0 + delta).view(self.layout.device_rows, 8)

    def reset_clock_words(self, clock: int) -> None:
        """After a resume: every version / clock word of this rank's block restarts at `clock`."""
        native.check(native.lib().psb_comm_set_clocks(self.block, self.F, self.world, int(clock), native._stream()),
                     "psb_comm_set_clocks")
        self.group_clock = [int(clock)] * self.groups
        torch.cuda.synchronize(self.device)

    def check(self) -> None:
        """Raises NativeError (PSB_E_TIMEOUT) if an in-kernel wait of this rank gave up; synchronises."""
        native.check(native.lib().psb_comm_check(self.block, self.F, self.world, native._stream()), "psb_comm_check")

    def status(self) -> int:
        ptr = native.lib().psb_comm_status(self.block, self.F, self.world)
        return int(_view(ptr, 1, self.device, torch.int32).item())

    def evaluate(self, weights: Optional[torch.Tensor] = None, impl: Optional[int] = None):
        """Weighted F1 / accuracy of the weights on the loaded test set (device finalise, 8-byte D2H)."""
        w = self.weights() if weights is None else weights
        torch.cuda.nvtx.range_push("psb.eval")
        native.evaluate(self.test_X, self.test_y, self.test_X.shape[0], w, self.F, self.layout.R, self.conf,
                        self.metrics_dev, self.impl if impl is None else impl)
        self.kernel_launches += 2
        self.metrics_host.copy_(self.metrics_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        torch.cuda.nvtx.range_pop()
        return float(self.metrics_host[0]), float(self.metrics_host[1])

    def wait_version(self, clock: Optional[int] = None) -> None:
        """Sequential mode, several ranks: enqueue a bounded device-side wait until every shard of `clock` (default:
        the clock after the last step) has reached this rank's replica.  A step only waits for the weights it pulls;
        whoever reads the replica right behind a step (evaluation, checkpoint) needs this first, because the last
        shards are written by whichever rank arrived last."""
        if self.comm_mode != native.COMM_SEQUENTIAL or self.world == 1:
            return
        native.check(native.lib().psb_comm_wait_version(self.block, self.F, self.world,
                                                        self.clock if clock is None else int(clock), self.spin_limit,
                                                        native._stream()), "psb_comm_wait_version")
        self.kernel_launches += 1

    def evaluate_workers(self, impl: Optional[int] = None):
        """Weighted F1 / accuracy of every worker's locally fitted model of the LAST step (weights it pulled + its
        delta; the worker log row, LogisticRegressionTaskSpark.java:186) — all W models in one launch, the sum formed
        while the weights are staged.  Returns (f1[W], accuracy[W]) as numpy arrays; synchronises."""
        nv, W = self.nv, self.W
        if not hasattr(self, "_wconf"):
            self._wconf = torch.zeros((W, 64), dtype=torch.int32, device=self.device)
            self._wmetrics = torch.zeros((W, 2), dtype=torch.float32, device=self.device)
            self._wmetrics_host = torch.zeros((W, 2), dtype=torch.float32).pin_memory()
        delta = self.scratch[:, 5 * nv: 6 * nv]              # solver scratch vector 5 = delta (csrc/worker_common.cuh)
        w0 = self.scratch[:, 6 * nv: 7 * nv]                 # vector 6 = the weights the step started from
        torch.cuda.nvtx.range_push("psb.eval_workers")
        native.evaluate_models(self.test_X, self.test_y, self.test_X.shape[0], w0, delta, self.scratch.shape[1], W,
                               self.F, self.layout.R, self._wconf, self._wmetrics,
                               self.impl if impl is None else impl)
        self.kernel_launches += 2
        self._wmetrics_host.copy_(self._wmetrics, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        torch.cuda.nvtx.range_pop()
        m = self._wmetrics_host.numpy()
        return m[:, 0].copy(), m[:, 1].copy()

    # ------------------------------------------------------------------ checkpoint / resume
    def state_dict(self) -> dict:
        """Everything this rank needs to continue: ring contents and windows, the buffer policy state (inter-arrival
        statistics), clocks and — for the relaxed models — this rank's authoritative shard.  Synchronises."""
        self._wait_slot_thread()
        torch.cuda.synchronize(self.device)
        L = native.lib()
        shard_f = ((self.nv // 4 + self.world - 1) // self.world) * 4
        shard = _view(L.psb_comm_shard_weights(self.block, self.F, self.world), shard_f, self.device)
        return {"rank": self.rank, "world": self.world, "W": self.W, "clock": self.clock,
                "group_clock": list(self.group_clock), "rows_ingested": self.rows_ingested,
                "X": self.X.cpu(), "labels": self.labels.cpu(), "win": self.win_host.clone(),
                "buffers": np.stack([b.get_state() for b in self.buffers]),
                "shard": shard.cpu(), "comm_mode": self.comm_mode,
                "worker_clocks": torch.from_numpy(self.worker_clocks().copy())}

    def load_state_dict(self, sd: dict, weights: Optional[np.ndarray] = None) -> None:
        """Inverse of state_dict (same rank / world / W).  `weights` [(F+1), 8]: the global weights of the
        checkpoint, written into both replicas (sequential mode); the relaxed modes restore the saved shard."""
        if (sd["rank"], sd["world"], sd["W"]) != (self.rank, self.world, self.W):
            raise ValueError("checkpoint was written by a different rank layout")
        self._wait_slot_thread()
        self.X.copy_(sd["X"]), self.labels.copy_(sd["labels"])
        self.win_host.copy_(sd["win"])
        for t in self._win:
            t.copy_(self.win_host)
        self._recompute_stats()
        for b, st in zip(self.buffers, sd["buffers"]):
            b.set_state(st)
        self._buf_handles = None
        self.rows_ingested = int(sd["rows_ingested"])
        self.clock = int(sd["clock"])
        L = native.lib()
        if weights is not None:
            w = torch.from_numpy(np.asarray(weights, dtype=np.float32).reshape(-1).copy()).to(self.device)
            for parity in (0, 1):
                _view(L.psb_comm_replica(self.block, self.F, self.world, parity), self.nv, self.device).copy_(w)
        shard = sd["shard"].to(self.device)
        _view(L.psb_comm_shard_weights(self.block, self.F, self.world), shard.numel(), self.device).copy_(shard)
        self.reset_clock_words(self.clock)
        if sd.get("group_clock") and len(sd["group_clock"]) == self.groups:
            self.group_clock = [int(c) for c in sd["group_clock"]]
        if self.comm_mode == native.COMM_WORKER and "worker_clocks" in sd:      # workers resume at their own clocks
            tab = sd["worker_clocks"].to(torch.int32).to(self.device)
            _view(L.psb_comm_worker_clocks(self.block, self.F, self.world), tab.numel(), self.device,
                  torch.int32).copy_(tab)
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        if self._slot_thread is not None:
            self._wait_slot_thread()
            self._slot_thread.shutdown(wait=True)
            self._slot_thread = None
        if self.block:
            torch.cuda.synchronize(self.device)
            for q, p in enumerate(self.peers):
                if q != self.rank and p and p != self.block and self._imported:
                    native.lib().psb_ipc_close(p)
            if self.owns_block:
                native.device_free(self.block)
            self.block = 0

    _imported = False


def _view(ptr: int, n: int, device, dtype=torch.float32) -> torch.Tensor:
    """Zero-copy torch view of raw device memory (through __cuda_array_interface__)."""
    class _Holder:
        pass
    h = _Holder()
    typestr = "<f4" if dtype == torch.float32 else "<i4"
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}
    return torch.as_tensor(h, device=device)


class VirtualCluster:
    """N ranks on ONE device, stepped in clock order (tests / single-GPU emulation of the protocol)."""

    def __init__(self, cfg: PSConfig, num_ranks: int, workers_per_rank: int, impl: int = native.IMPL_SIMT, w0=None,
                 **engine_kw):
        self.engines = [DeviceEngine(cfg, workers_per_rank, r, num_ranks, impl, w0=w0, **engine_kw)
                        for r in range(num_ranks)]
        peers = [e.block for e in self.engines]
        for e in self.engines:
            e.set_peers(peers)

    def step(self) -> None:
        for e in self.engines:
            e.step()

    def step_concurrent(self, num_clocks: int = 1) -> None:
        """Per-worker mode, eventual consistency: every virtual rank's launch on its own stream.  (Bounded delay makes
        the launches wait on one another; ranks that share a GPU are not guaranteed to run at the same time, so that
        model is exercised with one rank per GPU.)"""
        if self.engines[0].cfg.consistency_model > 0:
            raise native.NativeError("step_concurrent: bounded delay needs one rank per GPU")
        if not hasattr(self, "_streams"):
            self._streams = [[torch.cuda.Stream() for _ in range(e.groups)] for e in self.engines]
        cur = torch.cuda.current_stream()
        for e, sts in zip(self.engines, self._streams):
            for g, st in enumerate(sts):
                st.wait_stream(cur)
                e.step_group(g, num_clocks, stream=st)
        for sts in self._streams:
            for st in sts:
                cur.wait_stream(st)
        for e in self.engines:
            e.clock = min(e.group_clock)

    def close(self) -> None:
        for e in self.engines:
            e.close()
