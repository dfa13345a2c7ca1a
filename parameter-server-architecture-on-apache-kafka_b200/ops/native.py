"""ctypes binding of libpsb200.so (C ABI in ``include/psb200.h``).

The library is built in-tree by ``build.sh`` / ``__graft_entry__.build()`` for
sm_100a and loaded from this directory.  Host-only entry points (CSV tokenizer,
tracker, buffer policy) work on a machine without a GPU; anything that touches a
device raises ``NativeError`` with the library's own message when it fails —
there is no silent fallback to PyTorch ops.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PSB200_LIB") or os.path.join(_HERE, "libpsb200.so")   # override: experiment builds

CLASS_PAD = 8
MAX_RANKS = 8
HIST = 4
SCRATCH_VECS = 7 + 2 * HIST
SOLVER_LBFGS, SOLVER_SGD = 0, 1
COMM_NONE, COMM_SEQUENTIAL, COMM_ASYNC, COMM_WORKER = 0, 1, 2, 3
MAX_WORKERS = 4096
IMPL_SIMT, IMPL_TC = 0, 1


class NativeError(RuntimeError):
    pass


class PsbSolver(C.Structure):
    _fields_ = [("num_features", C.c_int), ("num_class_rows", C.c_int), ("cap", C.c_int), ("solver", C.c_int),
                ("max_iter", C.c_int), ("standardize", C.c_int), ("center", C.c_int), ("sgd_lr", C.c_float),
                ("inv_std", C.c_void_p)]


class PsbSparseComm(C.Structure):
    _fields_ = [("shard", C.c_void_p * MAX_RANKS), ("phase", C.c_void_p * MAX_RANKS), ("ticket", C.c_void_p),
                ("status", C.c_void_p), ("rank", C.c_int), ("num_ranks", C.c_int), ("rows_per_shard", C.c_int),
                ("spin_limit", C.c_ulonglong)]


class PsbSlabComm(C.Structure):
    _fields_ = [("base", PsbSparseComm), ("req", C.c_void_p * MAX_RANKS), ("req_cnt", C.c_void_p * MAX_RANKS),
                ("rep", C.c_void_p * MAX_RANKS), ("upd", C.c_void_p * MAX_RANKS), ("cnt", C.c_void_p),
                ("pos", C.c_void_p), ("cap", C.c_int), ("parity", C.c_int)]


class PsbComm(C.Structure):
    _fields_ = [("mode", C.c_int), ("rank", C.c_int), ("num_ranks", C.c_int), ("workers_per_rank", C.c_int),
                ("clock", C.c_int), ("max_delay", C.c_int), ("lr", C.c_float), ("spin_limit", C.c_ulonglong),
                ("peers", C.c_void_p * MAX_RANKS),
                # PSB_COMM_WORKER only (zero otherwise)
                ("num_clocks", C.c_int), ("worker_base", C.c_int), ("total_workers", C.c_int),
                ("delay_worker", C.c_int), ("delay_cycles", C.c_ulonglong), ("hist", C.c_void_p),
                ("home_remote", C.c_int), ("publish", C.c_int)]


class PsbClockRecord(C.Structure):
    _fields_ = [("loss", C.c_float), ("clock", C.c_int), ("lag", C.c_int), ("lead", C.c_int)]


CLOCK_RECORD_DTYPE = np.dtype([("loss", "<f4"), ("clock", "<i4"), ("lag", "<i4"), ("lead", "<i4")])


_lib: Optional[C.CDLL] = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not available():
        raise NativeError(f"{LIB_PATH} is missing: run ./build.sh (or __graft_entry__.build()) first")
    L = C.CDLL(LIB_PATH)
    vp, ip, fp = C.c_void_p, C.c_void_p, C.c_void_p
    L.psb_last_error.restype = C.c_char_p
    L.psb_version.restype = C.c_int
    L.psb_comm_block_bytes.restype = C.c_size_t
    L.psb_comm_block_bytes.argtypes = [C.c_int, C.c_int]
    L.psb_comm_init_block.argtypes = [vp, fp, C.c_int, C.c_int, C.c_int, vp]
    for name in ("psb_comm_replica",):
        getattr(L, name).restype = C.c_void_p
        getattr(L, name).argtypes = [vp, C.c_int, C.c_int, C.c_int]
    for name in ("psb_comm_shard_weights", "psb_comm_status", "psb_comm_version_words", "psb_comm_rank_clocks",
                 "psb_comm_worker_clocks", "psb_comm_published", "psb_comm_published_seq"):
        getattr(L, name).restype = C.c_void_p
        getattr(L, name).argtypes = [vp, C.c_int, C.c_int]
    L.psb_comm_set_clocks.argtypes = [vp, C.c_int, C.c_int, C.c_uint, vp]
    L.psb_worker_max_resident.argtypes = [C.POINTER(PsbSolver), C.c_int, C.c_int]
    L.psb_comm_check.argtypes = [vp, C.c_int, C.c_int, vp]
    L.psb_worker_scratch_bytes.restype = C.c_size_t
    L.psb_worker_scratch_bytes.argtypes = [C.c_int]
    L.psb_worker_step.argtypes = [C.POINTER(PsbSolver), C.POINTER(PsbComm), C.c_int, C.c_int, vp, ip, ip, fp, fp, fp,
                                  ip, fp, vp]
    L.psb_apply_delta.argtypes = [fp, fp, C.c_float, C.c_int, C.c_int, vp]
    L.psb_worker_step_csr.argtypes = [C.POINTER(PsbSolver), C.POINTER(PsbComm), C.c_int, C.c_int, ip, vp, ip, ip, ip, fp,
                                      fp, fp, ip, fp, vp]
    L.psb_ring_store_ell.argtypes = [vp, C.c_int, C.c_int, ip, vp, ip, ip, C.c_int, C.c_int, C.c_int, vp]
    L.psb_eval.argtypes = [vp, ip, C.c_int, fp, C.c_int, C.c_int, ip, fp, C.c_int, vp]
    L.psb_eval_finalize.argtypes = [ip, fp, vp]
    L.psb_sparse_step.argtypes = [C.POINTER(PsbSparseComm), C.c_int, C.c_uint, C.c_uint, vp, vp, vp, vp, C.c_int, C.c_int,
                                  C.c_int, C.c_float, vp, vp, vp]
    L.psb_sparse_eval.argtypes = [C.POINTER(PsbSparseComm), vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp, vp]
    L.psb_sparse_step_slab.argtypes = [C.POINTER(PsbSlabComm), C.c_int, vp, vp, vp, vp, C.c_int, C.c_longlong, C.c_int,
                                       C.c_int, C.c_float, vp, C.c_int, vp]
    L.psb_ring_scatter.argtypes = [vp, C.c_int, C.c_int, vp, ip, C.c_int, C.c_int, vp]
    L.psb_ring_store_dense.argtypes = [vp, C.c_int, ip, ip, C.c_int, vp, ip, C.c_int, C.c_int, vp]
    L.psb_malloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
    L.psb_free.argtypes = [vp]
    L.psb_csv_count.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                                C.POINTER(C.c_int)]
    L.psb_ring_store_bitmap.argtypes = [vp, ip, vp, ip, ip, C.c_int, vp, ip, C.c_int, C.c_int, vp]
    L.psb_pack_bitmap_csr.restype = C.c_longlong
    L.psb_pack_bitmap_csr.argtypes = [vp, vp, vp, C.c_longlong, C.c_int, vp, vp, vp]
    L.psb_pack_bitmap_dense.restype = C.c_longlong
    L.psb_pack_bitmap_dense.argtypes = [vp, C.c_longlong, C.c_int, vp, vp, vp]
    L.psb_h2d_copies.argtypes = [C.c_int, vp, vp, vp, vp]
    L.psb_window_stats_update.argtypes = [vp, ip, C.c_int, C.c_int, C.c_int, vp, fp, vp]
    L.psb_d2h_copy.argtypes = [vp, vp, C.c_size_t, vp]
    L.psb_d2d_copy.argtypes = [vp, vp, C.c_size_t, vp]
    L.psb_pack_palette_dense.restype = C.c_longlong
    L.psb_pack_palette_dense.argtypes = [vp, C.c_longlong, C.c_int, vp, vp, vp, C.c_longlong]
    L.psb_ring_store_palette.argtypes = [vp, ip, vp, ip, ip, C.c_int, vp, ip, C.c_int, C.c_int, vp]
    L.psb_csv_parse.argtypes = [C.c_char_p, C.c_int, vp, vp, vp, vp, C.c_longlong, C.c_longlong]
    L.psb_csv_open.restype = C.c_void_p
    L.psb_csv_open.argtypes = [C.c_char_p, C.c_int, C.c_int]
    L.psb_csv_shape.argtypes = [vp, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_int)]
    L.psb_csv_fill.argtypes = [vp, vp, vp, vp, vp]
    L.psb_csv_close.argtypes = [vp]
    L.psb_csv_close.restype = None
    L.psb_tracker_create.restype = C.c_void_p
    L.psb_tracker_create.argtypes = [C.c_int, C.c_int]
    L.psb_tracker_destroy.argtypes = [vp]
    L.psb_tracker_on_gradient.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int]
    L.psb_tracker_clocks.argtypes = [vp, vp]
    L.psb_buffer_create.restype = C.c_void_p
    L.psb_buffer_create.argtypes = [C.c_int, C.c_int, C.c_float, C.c_int]
    L.psb_buffer_destroy.argtypes = [vp]
    L.psb_buffer_on_event.argtypes = [vp, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.psb_buffer_inserted.restype = C.c_longlong
    L.psb_buffer_inserted.argtypes = [vp]
    L.psb_buffers_on_batch.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp, vp]
    L.psb_buffers_on_batch2.argtypes = [vp, C.c_int, vp, vp, C.c_int, vp, vp, vp, C.POINTER(C.c_int)]
    L.psb_buffer_get_state.argtypes = [vp, vp, C.c_int]
    L.psb_buffer_set_ring.argtypes = [vp, C.c_int]
    L.psb_buffer_set_state.argtypes = [vp, vp, C.c_int]
    L.psb_comm_wait_version.argtypes = [vp, C.c_int, C.c_int, C.c_uint, C.c_ulonglong, vp]
    L.psb_eval_models.argtypes = [vp, ip, C.c_int, fp, fp, C.c_size_t, C.c_int, C.c_int, C.c_int, ip, fp, C.c_int, vp]
    L.psb_producer_start.restype = C.c_void_p
    L.psb_producer_start.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     vp, vp, vp, vp, C.c_int]
    L.psb_producer_start_packed.restype = C.c_void_p
    L.psb_producer_start_packed.argtypes = L.psb_producer_start.argtypes + [vp, vp, vp]
    L.psb_producer_start_palette.restype = C.c_void_p
    L.psb_producer_start_palette.argtypes = L.psb_producer_start.argtypes + [vp, vp, vp, C.c_longlong]
    L.psb_producer_poll.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.psb_producer_release.argtypes = [vp, C.c_int]
    L.psb_producer_rows_sent.restype = C.c_longlong
    L.psb_producer_rows_sent.argtypes = [vp]
    L.psb_producer_stop.argtypes = [vp]
    L.psb_ipc_export.argtypes = [vp, vp]
    L.psb_ipc_import.argtypes = [vp, C.POINTER(C.c_void_p)]
    L.psb_ipc_close.argtypes = [vp]
    L.psb_enable_peer_access.argtypes = [C.c_int]
    L.psb_set_l2_fetch_granularity.argtypes = [C.c_int]
    L.psb_static_scratch_bytes.restype = C.c_size_t
    L.psb_static_scratch_bytes.argtypes = []
    L.psb_worker_step_static.argtypes = [C.POINTER(PsbSolver), C.POINTER(PsbComm), C.c_int, C.c_int, vp, ip, ip, fp, fp,
                                         fp, ip, vp]
    _lib = L
    return L


def check(rc: int, what: str = "") -> int:
    if rc < 0:
        raise NativeError(f"{what}: {lib().psb_last_error().decode()} (code {rc})")
    return rc


def _ptr(t) -> int:
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if isinstance(t, int):                      # raw device address (memory mapped from another process)
        return t
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


def _stream(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


# ------------------------------------------------------------------ host runtime wrappers
def csv_to_csr(path: str, has_header: bool = True, threads: int = 0):
    """The whole CSV as a CSR batch: one pass over the text, tokenised by `threads` threads (0: up to 8)."""
    from ..data.csr import CSRBatch
    L = lib()
    h = L.psb_csv_open(path.encode(), int(has_header), int(threads))
    if not h:
        raise NativeError("csv_open: " + L.psb_last_error().decode())
    try:
        n_rows, n_nnz, n_cols = C.c_longlong(), C.c_longlong(), C.c_int()
        check(L.psb_csv_shape(h, C.byref(n_rows), C.byref(n_nnz), C.byref(n_cols)), "csv_shape")
        indptr = np.zeros(n_rows.value + 1, dtype=np.int64)
        indices = np.zeros(max(1, n_nnz.value), dtype=np.int32)
        values = np.zeros(max(1, n_nnz.value), dtype=np.float32)
        labels = np.zeros(max(1, n_rows.value), dtype=np.int32)
        check(L.psb_csv_fill(h, indptr.ctypes.data, indices.ctypes.data, values.ctypes.data, labels.ctypes.data), "csv_fill")
    finally:
        L.psb_csv_close(h)
    return CSRBatch(indptr, indices[: n_nnz.value], values[: n_nnz.value], labels[: n_rows.value], n_cols.value)


def csv_to_csr_two_pass(path: str, has_header: bool = True):
    """The same through psb_csv_count + psb_csv_parse (no allocation inside the library: the caller owns every array)."""
    from ..data.csr import CSRBatch
    L = lib()
    n_rows, n_nnz, n_cols = C.c_longlong(), C.c_longlong(), C.c_int()
    check(L.psb_csv_count(path.encode(), int(has_header), C.byref(n_rows), C.byref(n_nnz), C.byref(n_cols)), "csv_count")
    indptr = np.zeros(n_rows.value + 1, dtype=np.int64)
    indices = np.zeros(max(1, n_nnz.value), dtype=np.int32)
    values = np.zeros(max(1, n_nnz.value), dtype=np.float32)
    labels = np.zeros(max(1, n_rows.value), dtype=np.int32)
    check(L.psb_csv_parse(path.encode(), int(has_header), indptr.ctypes.data, indices.ctypes.data, values.ctypes.data,
                          labels.ctypes.data, n_rows.value, n_nnz.value), "csv_parse")
    return CSRBatch(indptr, indices[: n_nnz.value], values[: n_nnz.value], labels[: n_rows.value], n_cols.value)


class PackedRows:
    """Packed rows in pinned host memory: the sparse wire format of ingestion (LabeledData.java:14-28 without the
    keys).  Lossless w.r.t. the fp16 rings.  Two encodings of the non-zeros, both behind a presence bitmap:
      * ``palette=False``: the fp16 values in feature order (128 B + 2 B per non-zero at F = 1024);
      * ``palette=True`` : per row the <= 16 distinct fp16 values it uses + a 4-bit code per non-zero (hashed
        bag-of-words rows hold a handful of distinct values: count / norm, signed); rows with more distinct
        values fall back to raw fp16 values.  128 B + ~0.5 B per non-zero."""

    def __init__(self, max_rows: int, num_features: int, pin: bool = True, palette: bool = False):
        import torch
        assert num_features % 32 == 0, "bitmap packing needs num_features % 32 == 0"
        self.max_rows, self.F, self.n, self.nnz, self.palette = max_rows, num_features, 0, 0, bool(palette)
        mk = (lambda *a, **k: torch.zeros(*a, **k).pin_memory()) if pin and torch.cuda.is_available() else torch.zeros
        self.bitmap = mk((max_rows, num_features // 32), dtype=torch.int32)
        self.row_off = mk(max_rows + 1, dtype=torch.int32)
        if self.palette:
            self.payload = mk(max_rows * (2 * num_features + 2), dtype=torch.uint8)
            self.payload_bytes = 0
        else:
            self.values = mk(max_rows * num_features, dtype=torch.float16)
        self.labels = mk(max_rows, dtype=torch.int32)

    def pack_dense(self, rows_f16, labels) -> "PackedRows":
        """rows_f16: torch/numpy fp16 [n, F] on the host."""
        n = int(rows_f16.shape[0])
        assert n <= self.max_rows and int(rows_f16.shape[1]) == self.F
        if self.palette:
            self.payload_bytes = int(check(lib().psb_pack_palette_dense(
                _ptr(rows_f16), n, self.F, _ptr(self.bitmap), _ptr(self.row_off), _ptr(self.payload),
                self.payload.numel()), "pack_palette_dense"))
            self.nnz = -1
        else:
            self.nnz = int(check(lib().psb_pack_bitmap_dense(_ptr(rows_f16), n, self.F, _ptr(self.bitmap),
                                                             _ptr(self.row_off), _ptr(self.values)), "pack_bitmap_dense"))
        self.n = n
        self.labels[:n] = _as_int32_tensor(labels)[:n]
        return self

    def pack_csr(self, batch) -> "PackedRows":
        n = len(batch)
        assert n <= self.max_rows and batch.num_features <= self.F
        if self.palette:
            dense = np.zeros((n, self.F), dtype=np.float16)
            dense[:, :batch.num_features] = batch.to_dense(np.float16)
            return self.pack_dense(dense, batch.labels)
        indptr = np.ascontiguousarray(batch.indptr, dtype=np.int64)
        indices = np.ascontiguousarray(batch.indices, dtype=np.int32)
        values = np.ascontiguousarray(batch.values, dtype=np.float32)
        self.nnz = int(check(lib().psb_pack_bitmap_csr(indptr.ctypes.data, indices.ctypes.data, values.ctypes.data, n,
                                                       self.F, _ptr(self.bitmap), _ptr(self.row_off),
                                                       _ptr(self.values)), "pack_bitmap_csr"))
        self.n = n
        self.labels[:n] = _as_int32_tensor(batch.labels)[:n]
        return self

    @property
    def nbytes(self) -> int:
        """Bytes that cross PCIe for this batch (bitmap + offsets + values or payload + labels)."""
        body = self.payload_bytes if self.palette else self.nnz * 2
        return self.n * (self.F // 8) + (self.n + 1) * 4 + body + self.n * 4

    def to_dense(self) -> np.ndarray:
        """Reference expansion on the host (tests)."""
        out = np.zeros((self.n, self.F), dtype=np.float16)
        bits = np.unpackbits(self.bitmap[: self.n].numpy().view(np.uint8), axis=1, bitorder="little").astype(bool)
        if not self.palette:
            out[bits] = self.values[: self.nnz].numpy()
            return out
        pay, off = self.payload.numpy(), self.row_off.numpy()
        for r in range(self.n):
            p = pay[off[r]: off[r + 1]]
            npal, nnz = int(p[:2].view(np.uint16)[0]), int(bits[r].sum())
            if npal == 0:
                vals = p[2: 2 + 2 * nnz].view(np.float16)
            else:
                pal = p[2: 2 + 2 * npal].view(np.float16)
                c = p[2 + 2 * npal: 2 + 2 * npal + (nnz + 1) // 2]
                codes = np.stack([c & 15, c >> 4], axis=1).reshape(-1)[:nnz]
                vals = pal[codes]
            out[r, bits[r]] = vals
        return out


def _as_int32_tensor(x):
    import torch
    if isinstance(x, np.ndarray):
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.int32))
    return x.to(torch.int32)


def ring_store_bitmap(bitmap, row_off, values, labels_in, worker_slot, n: int, X, labels, cap: int, num_features: int,
                      stream=None) -> None:
    check(lib().psb_ring_store_bitmap(_ptr(bitmap), _ptr(row_off), _ptr(values), _ptr(labels_in), _ptr(worker_slot), n,
                                      _ptr(X), _ptr(labels), cap, num_features, _stream(stream)),
          "psb_ring_store_bitmap")


def h2d_copies(copies, stream=None) -> None:
    """copies: iterable of (device tensor, pinned host tensor, nbytes); one native call, asynchronous on `stream`."""
    copies = [c for c in copies if c[2] > 0]
    n = len(copies)
    dst = (C.c_void_p * n)(*[c[0].data_ptr() for c in copies])
    src = (C.c_void_p * n)(*[c[1].data_ptr() for c in copies])
    nb = (C.c_size_t * n)(*[int(c[2]) for c in copies])
    check(lib().psb_h2d_copies(n, dst, src, nb, _stream(stream)), "psb_h2d_copies")


def ring_store_palette(bitmap, row_off, payload, labels_in, worker_slot, n: int, X, labels, cap: int,
                       num_features: int, stream=None) -> None:
    check(lib().psb_ring_store_palette(_ptr(bitmap), _ptr(row_off), _ptr(payload), _ptr(labels_in), _ptr(worker_slot), n,
                                       _ptr(X), _ptr(labels), cap, num_features, _stream(stream)),
          "psb_ring_store_palette")


class NativeTracker:
    def __init__(self, num_workers: int, consistency_model: int):
        self.n = num_workers
        self.h = lib().psb_tracker_create(num_workers, consistency_model)
        if not self.h:
            raise NativeError("psb_tracker_create failed")
        self._buf = np.zeros(2 * num_workers, dtype=np.int32)

    def on_gradient(self, worker: int, clock: int):
        n = check(lib().psb_tracker_on_gradient(self.h, worker, clock, self._buf.ctypes.data, self.n), "tracker")
        return [(int(self._buf[2 * i]), int(self._buf[2 * i + 1])) for i in range(n)]

    def clocks(self):
        out = np.zeros(self.n, dtype=np.int32)
        lib().psb_tracker_clocks(self.h, out.ctypes.data)
        return out.tolist()

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.psb_tracker_destroy(self.h)
            self.h = None


class NativeBuffer:
    def __init__(self, min_size: int, max_size: int, bc: float, window: int = 500, ring: int = 0):
        """ring: physical slots of the device ring when it holds more than max_size (spare slots)."""
        self.h = lib().psb_buffer_create(min_size, max_size, bc, window)
        if not self.h:
            raise NativeError("psb_buffer_create failed")
        self.window = window if window > 0 else 500
        if ring and ring != max_size:
            check(lib().psb_buffer_set_ring(self.h, ring), "psb_buffer_set_ring")
        self.head = self.count = self.target = 0

    def ring_total(self) -> int:
        """rows ever inserted (numTuplesSeen of the reference worker log)"""
        return int(lib().psb_buffer_inserted(self.h))

    def on_event(self, now_ms: float) -> int:
        h, c, t = C.c_int(), C.c_int(), C.c_int()
        slot = lib().psb_buffer_on_event(self.h, float(now_ms), C.byref(h), C.byref(c), C.byref(t))
        self.head, self.count, self.target = h.value, c.value, t.value
        return slot

    def get_state(self) -> np.ndarray:
        """Policy state (inter-arrival window + ring cursors) as a float64 vector, for checkpoints."""
        out = np.zeros(9 + self.window, dtype=np.float64)
        n = lib().psb_buffer_get_state(self.h, out.ctypes.data, len(out))
        if n != len(out):
            raise NativeError("psb_buffer_get_state failed")
        return out

    def set_state(self, state: np.ndarray) -> None:
        state = np.ascontiguousarray(state, dtype=np.float64)
        check(lib().psb_buffer_set_state(self.h, state.ctypes.data, len(state)), "psb_buffer_set_state")
        self.head, self.count = int(state[6]), int(state[7])

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.psb_buffer_destroy(self.h)
            self.h = None


def buffer_handles(buffers):
    """ctypes array of the buffers' handles (build once per buffer set, pass as ``handles=``)."""
    return (C.c_void_p * len(buffers))(*[b.h for b in buffers])


def buffers_on_batch(buffers, workers: np.ndarray, now_ms: np.ndarray, worker_slot: np.ndarray, win: np.ndarray,
                     handles=None, change: Optional[np.ndarray] = None) -> int:
    """buffers: list of NativeBuffer; fills worker_slot[n,2] / win[W,2] (int32 numpy views, may be pinned) and, when
    given, change[W,4] = (head before, count before, rows added, count after).  Returns the most rows any worker got."""
    n = len(workers)
    arr = handles if handles is not None else buffer_handles(buffers)
    workers = np.ascontiguousarray(workers, dtype=np.int32)
    now_ms = np.ascontiguousarray(now_ms, dtype=np.float64)
    most = C.c_int(0)
    check(lib().psb_buffers_on_batch2(arr, len(buffers), workers.ctypes.data, now_ms.ctypes.data, n,
                                      worker_slot.ctypes.data, win.ctypes.data,
                                      change.ctypes.data if change is not None else None, C.byref(most)),
          "psb_buffers_on_batch")
    return most.value


class StreamProducer:
    """Host producer thread filling a ring of pinned batches (see psb_producer_start)."""

    def __init__(self, path: str, num_workers_total: int, worker_begin: int, worker_count: int, num_features: int,
                 p_ms: int, batch_rows: int = 1024, num_batches: int = 4, has_header: bool = True, flush_ms: int = 50,
                 packed=False):
        """packed: False (dense fp16 batches only), True (also bitmap + fp16 values) or "palette" (also bitmap +
        per-row palette + 4-bit codes)."""
        import torch
        pin = torch.cuda.is_available()
        mk = (lambda *s, dtype: torch.zeros(*s, dtype=dtype).pin_memory()) if pin else \
            (lambda *s, dtype: torch.zeros(*s, dtype=dtype))
        self.rows = mk(num_batches, batch_rows, num_features, dtype=torch.float16)
        self.labels = mk(num_batches, batch_rows, dtype=torch.int32)
        self.workers = mk(num_batches, batch_rows, dtype=torch.int32)
        self.times = mk(num_batches, batch_rows, dtype=torch.float64)
        self.is_packed = bool(packed) and num_features % 32 == 0
        self.is_palette = self.is_packed and packed == "palette"
        self.F, self.batch_rows = num_features, batch_rows
        common = (path.encode(), int(has_header), num_workers_total, worker_begin, worker_count, num_features, int(p_ms),
                  batch_rows, num_batches, self.rows.data_ptr(), self.labels.data_ptr(), self.workers.data_ptr(),
                  self.times.data_ptr(), flush_ms)
        if self.is_packed:
            self.bitmap = mk(num_batches, batch_rows, num_features // 32, dtype=torch.int32)
            self.row_off = mk(num_batches, batch_rows + 1, dtype=torch.int32)
        if self.is_palette:
            stride = batch_rows * (2 * num_features + 2)
            self.payload = mk(num_batches, stride, dtype=torch.uint8)
            self.h = lib().psb_producer_start_palette(*common, self.bitmap.data_ptr(), self.row_off.data_ptr(),
                                                      self.payload.data_ptr(), stride)
        elif self.is_packed:
            self.values = mk(num_batches, batch_rows * num_features, dtype=torch.float16)
            self.h = lib().psb_producer_start_packed(*common, self.bitmap.data_ptr(), self.row_off.data_ptr(),
                                                     self.values.data_ptr())
        else:
            self.h = lib().psb_producer_start_packed(*common, None, None, None)
        if not self.h:
            raise NativeError("psb_producer_start: " + lib().psb_last_error().decode())

    def poll(self):
        """(batch, n, eof): batch is None when nothing is ready; eof is 0, 1 (done) or 2 (file not found)."""
        b, n, e = C.c_int(), C.c_int(), C.c_int()
        if lib().psb_producer_poll(self.h, C.byref(b), C.byref(n), C.byref(e)):
            return b.value, n.value, 0
        if e.value == 3:
            raise NativeError("the rows file is damaged or was written for another number of features")
        return None, 0, e.value

    def packed(self, batch: int, n: int) -> "PackedRows":
        """The bitmap-packed form of ready batch `batch` (views of the producer's pinned buffers, no copy)."""
        assert self.is_packed, "StreamProducer(packed=True) required"
        pk = PackedRows.__new__(PackedRows)
        pk.max_rows, pk.F, pk.n, pk.palette = self.batch_rows, self.F, n, self.is_palette
        pk.bitmap, pk.row_off, pk.labels = self.bitmap[batch], self.row_off[batch], self.labels[batch]
        if self.is_palette:
            pk.payload, pk.payload_bytes, pk.nnz = self.payload[batch], int(self.row_off[batch, n]), -1
        else:
            pk.values, pk.nnz = self.values[batch], int(self.row_off[batch, n])
        return pk

    def release(self, batch: int) -> None:
        lib().psb_producer_release(self.h, batch)

    def rows_sent(self) -> int:
        return lib().psb_producer_rows_sent(self.h)

    def stop(self) -> None:
        if self.h:
            lib().psb_producer_stop(self.h)
            self.h = None

    def __del__(self):
        try:
            self.stop()
        except Exception:
            pass


# ------------------------------------------------------------------ device wrappers
def make_solver(num_features, num_class_rows, cap, solver="lbfgs2", max_iter=2, standardize=True, center=True,
                sgd_lr=0.5) -> PsbSolver:
    return PsbSolver(num_features, num_class_rows, cap, SOLVER_LBFGS if solver == "lbfgs2" else SOLVER_SGD, max_iter,
                     int(standardize), int(center), sgd_lr)


def comm_none() -> PsbComm:
    c = PsbComm()
    c.mode = COMM_NONE
    c.num_ranks = 1
    return c


def worker_step(solver: PsbSolver, comm: PsbComm, num_workers: int, impl: int, X, labels, win, w_in, delta, loss,
                info, scratch, stream=None) -> None:
    check(lib().psb_worker_step(C.byref(solver), C.byref(comm), num_workers, impl, _ptr(X), _ptr(labels), _ptr(win),
                                _ptr(w_in), _ptr(delta), _ptr(loss), _ptr(info), _ptr(scratch), _stream(stream)),
          "psb_worker_step")


def apply_delta(w, delta, lr: float, num_features: int, skip_class: int = -1, stream=None) -> None:
    check(lib().psb_apply_delta(_ptr(w), _ptr(delta), lr, num_features, skip_class, _stream(stream)), "psb_apply_delta")


def evaluate(X, labels, n: int, w, num_features: int, num_class_rows: int, conf, out, impl: int = 0, stream=None):
    check(lib().psb_eval(_ptr(X), _ptr(labels), n, _ptr(w), num_features, num_class_rows, _ptr(conf), _ptr(out), impl,
                         _stream(stream)), "psb_eval")


def evaluate_models(X, labels, n: int, w, w_add, model_stride: int, num_models: int, num_features: int,
                    num_class_rows: int, conf, out, impl: int = 0, stream=None):
    """Weighted F1 / accuracy of `num_models` weight sets in one launch (conf int32[M, 64], out float32[M, 2])."""
    check(lib().psb_eval_models(_ptr(X), _ptr(labels), n, _ptr(w), _ptr(w_add), model_stride, num_models, num_features,
                                num_class_rows, _ptr(conf), _ptr(out), impl, _stream(stream)), "psb_eval_models")


def window_stats_update(X, change, num_workers: int, cap: int, num_features: int, sums, inv_std, stream=None) -> None:
    check(lib().psb_window_stats_update(_ptr(X), _ptr(change), num_workers, cap, num_features, _ptr(sums),
                                        _ptr(inv_std), _stream(stream)), "psb_window_stats_update")


def ring_scatter(stage, n: int, stride_words: int, X, labels, cap: int, num_features: int, stream=None) -> None:
    check(lib().psb_ring_scatter(_ptr(stage), n, stride_words, _ptr(X), _ptr(labels), cap, num_features,
                                 _stream(stream)), "psb_ring_scatter")


def ring_store_dense(rows, labels_in, worker_slot, n: int, X, labels, cap: int, num_features: int, stream=None):
    import torch
    src_half = int(rows.dtype == torch.float16)
    check(lib().psb_ring_store_dense(_ptr(rows), src_half, _ptr(labels_in), _ptr(worker_slot), n, _ptr(X),
                                     _ptr(labels), cap, num_features, _stream(stream)), "psb_ring_store_dense")


def device_malloc(nbytes: int) -> int:
    p = C.c_void_p()
    check(lib().psb_malloc(C.byref(p), nbytes), "psb_malloc")
    return p.value


def device_free(ptr: int) -> None:
    check(lib().psb_free(ptr), "psb_free")


def ipc_export(ptr: int) -> bytes:
    buf = (C.c_ubyte * 64)()
    check(lib().psb_ipc_export(ptr, buf), "psb_ipc_export")
    return bytes(buf)


def ipc_import(handle: bytes) -> int:
    buf = (C.c_ubyte * 64).from_buffer_copy(handle)
    p = C.c_void_p()
    check(lib().psb_ipc_import(buf, C.byref(p)), "psb_ipc_import")
    return p.value


def scratch_floats(num_features: int) -> int:
    return lib().psb_worker_scratch_bytes(num_features) // 4


def comm_block_bytes(num_features: int, num_ranks: int) -> int:
    return lib().psb_comm_block_bytes(num_features, num_ranks)
