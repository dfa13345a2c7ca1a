"""Peer-memory transport between the ServerAppRunner and WorkerAppRunner processes.

The reference's two JVMs meet at a Kafka broker (``ServerAppRunner.java:63``, ``ServerApp.java:31-42``: the server
creates the three topics, ``BaseKafkaApp.java:27-33``).  Here the same address is a ``TCPStore`` rendezvous that is
used ONCE, to hand over a small JSON description of the run and two CUDA IPC handles; afterwards nothing but
device memory connects the two processes:

* ``WEIGHTS_TOPIC`` / ``GRADIENTS_TOPIC`` -> the server's peer-memory block (``csrc/common.cuh: CommLayout``),
  mapped into the worker process; the worker kernels pull, push and apply in it (``csrc/comm.cuh``) and the server
  reads clocks and published weights out of it;
* ``INPUT_DATA`` -> ``InputTopic``: a ring of ``num_sets`` batch slots in device memory plus a control page.  The
  server's producer side copies packed rows (presence bitmap + palette payload, ``psb_pack_palette_dense``) from pinned
  host memory straight into a free slot and then bumps ``produced``; the worker stores the slot's rows into its input
  rings with ``psb_ring_store_palette`` (reading the slot in place) and then bumps ``consumed``.  Both counters are
  written with stream-ordered 4-byte copies, so a counter never overtakes the data it announces.

The control page is plain memory with single-writer words, which makes the protocol testable without a GPU: the
``HostRegion`` below keeps the same bytes in ``multiprocessing.shared_memory``.
"""
from __future__ import annotations

import json
import time
from dataclasses import dataclass
from typing import Optional

import numpy as np

# control words (int32 index in the control page); S = written by the server app, W = by the worker app
CTL_PRODUCED = 0        # S: batches published so far
CTL_EOF = 1             # S: 1 once the producer has published its last batch
CTL_STOP = 2            # S: 1 = stop training now
CTL_SERVER_LOGGED = 3   # S: clocks whose weights the server has scored and logged
CTL_CONSUMED = 16       # W: batches stored into the rings so far (own 64-byte line)
CTL_WORKER_DONE = 17    # W: 1 once the worker app has run its last clock
CTL_WORKER_CLOCK = 18   # W: clocks completed (host-side mirror, for the server's progress watchdog)
CTL_WORDS = 64

HDR_N, HDR_PAYLOAD_BYTES, HDR_SEQ = 0, 1, 2
HDR_WORDS = 16


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


@dataclass
class InputTopicLayout:
    """Byte layout of the INPUT_DATA topic: control page, then ``num_sets`` identical batch slots."""
    num_features: int
    rows_per_set: int
    num_sets: int = 4

    def __post_init__(self):
        F, R = self.num_features, self.rows_per_set
        if F % 32:
            raise ValueError("the packed row format needs num_features % 32 == 0")
        o = 0
        self.off_hdr = o; o = _align(o + HDR_WORDS * 4)
        self.off_workers = o; o = _align(o + R * 4)
        self.off_times = o; o = _align(o + R * 8)
        self.off_labels = o; o = _align(o + R * 4)
        self.off_bitmap = o; o = _align(o + R * (F // 32) * 4)
        self.off_row_off = o; o = _align(o + (R + 1) * 4)
        self.off_payload = o; o = _align(o + R * (2 * F + 2))
        self.set_bytes = o
        self.payload_capacity = R * (2 * F + 2)
        self.off_sets = _align(CTL_WORDS * 4)
        self.bytes = self.off_sets + self.num_sets * self.set_bytes

    def set_base(self, j: int) -> int:
        return self.off_sets + (j % self.num_sets) * self.set_bytes

    def describe(self) -> dict:
        return {"num_features": self.num_features, "rows_per_set": self.rows_per_set, "num_sets": self.num_sets}


class HostRegion:
    """The topic's bytes in host shared memory (CPU tests, and the statement of what a region must do)."""

    def __init__(self, nbytes: int, name: Optional[str] = None, create: bool = True):
        from multiprocessing import shared_memory
        self.shm = shared_memory.SharedMemory(name=name, create=create, size=nbytes if create else 0)
        self.buf = np.frombuffer(self.shm.buf, dtype=np.uint8, count=nbytes)
        if create:
            self.buf[:] = 0
        self.nbytes, self.owner = nbytes, create

    @property
    def name(self) -> str:
        return self.shm.name

    def write(self, offset: int, src: np.ndarray, stream=None) -> None:
        raw = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
        self.buf[offset: offset + raw.size] = raw

    def read(self, offset: int, nbytes: int, stream=None) -> np.ndarray:
        return self.buf[offset: offset + nbytes].copy()

    def address(self, offset: int) -> int:
        return self.buf.ctypes.data + offset

    def close(self) -> None:
        self.buf = None
        self.shm.close()
        if self.owner:
            try:
                self.shm.unlink()
            except FileNotFoundError:
                pass


class DeviceRegion:
    """The topic's bytes in device memory: a raw allocation of this process, or another process' allocation mapped
    through CUDA IPC.  Writes are asynchronous copies from pinned memory on ``stream``; reads synchronise."""

    def __init__(self, nbytes: int, ptr: Optional[int] = None):
        import torch
        from ..ops import native
        self.native, self.torch = native, torch
        self.nbytes, self.owner = nbytes, ptr is None
        self.ptr = native.device_malloc(nbytes) if ptr is None else int(ptr)
        self.stream = torch.cuda.Stream()
        self._pin_w = [torch.zeros(64, dtype=torch.uint8).pin_memory() for _ in range(16)]   # small-word writes
        self._pin_ev = [None] * 16                     # copy out of pinned word i has completed
        self._pin_i = 0
        self._pin_r = torch.zeros(1 << 16, dtype=torch.uint8).pin_memory()
        if self.owner:
            from ..engine.device_engine import _view
            _view(self.ptr, nbytes // 4, torch.device("cuda", torch.cuda.current_device()), torch.int32).zero_()
            torch.cuda.synchronize()

    def handle(self) -> bytes:
        return self.native.ipc_export(self.ptr)

    @staticmethod
    def from_handle(handle: bytes, nbytes: int) -> "DeviceRegion":
        from ..ops import native
        r = DeviceRegion(nbytes, ptr=native.ipc_import(handle))
        r.imported = True
        return r

    imported = False

    def write(self, offset: int, src, stream=None) -> None:
        """src: a pinned torch tensor (asynchronous) or a small numpy array (staged through a pinned word)."""
        torch = self.torch
        st = stream or self.stream
        if isinstance(src, np.ndarray):
            raw = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
            assert raw.size <= 64, "numpy writes are for control words; pass pinned tensors for bulk data"
            i = self._pin_i
            self._pin_i = (i + 1) % len(self._pin_w)
            if self._pin_ev[i] is not None:
                self._pin_ev[i].synchronize()           # the word's previous copy has left the pinned buffer
            pin = self._pin_w[i]
            pin[: raw.size] = torch.from_numpy(raw.copy())
            self.native.h2d_copies([(_Addr(self.ptr + offset), pin, raw.size)], stream=st)
            self._pin_ev[i] = torch.cuda.Event()
            self._pin_ev[i].record(st)
            return
        self.native.h2d_copies([(_Addr(self.ptr + offset), src, src.numel() * src.element_size())], stream=st)

    def write_bytes(self, offset: int, src, nbytes: int, stream=None) -> None:
        self.native.h2d_copies([(_Addr(self.ptr + offset), src, nbytes)], stream=stream or self.stream)

    def read(self, offset: int, nbytes: int, stream=None) -> np.ndarray:
        st = stream or self.stream
        if nbytes > self._pin_r.numel():
            self._pin_r = self.torch.zeros(nbytes, dtype=self.torch.uint8).pin_memory()
        self.native.check(self.native.lib().psb_d2h_copy(self._pin_r.data_ptr(), self.ptr + offset, nbytes,
                                                         st.cuda_stream), "psb_d2h_copy")
        st.synchronize()
        return self._pin_r[:nbytes].numpy().copy()

    def address(self, offset: int) -> int:
        return self.ptr + offset

    def close(self) -> None:
        if self.ptr:
            self.torch.cuda.synchronize()
            if self.owner:
                self.native.device_free(self.ptr)
            elif self.imported:
                self.native.lib().psb_ipc_close(self.ptr)
            self.ptr = 0


class _Addr:
    """Duck-typed 'tensor' for native.h2d_copies: a raw device address."""

    def __init__(self, ptr: int):
        self.ptr = ptr

    def data_ptr(self) -> int:
        return self.ptr


class InputTopicWriter:
    """Server side of the INPUT_DATA topic (``CsvProducer.runProducer`` -> ``producer.send``, CsvProducer.java:61-65)."""

    def __init__(self, region, layout: InputTopicLayout):
        self.r, self.L = region, layout
        self.produced = 0
        self.consumed = 0

    def _ctl(self, word: int) -> int:
        return int(self.r.read(word * 4, 4).view(np.int32)[0])

    def has_room(self) -> bool:
        if self.produced - self.consumed >= self.L.num_sets:
            self.consumed = self._ctl(CTL_CONSUMED)
        return self.produced - self.consumed < self.L.num_sets

    def publish(self, n: int, workers, times, labels, bitmap, row_off, payload, payload_bytes: int) -> bool:
        """One batch of packed rows (host arrays / pinned tensors).  False when every slot is still unread."""
        if not self.has_room():
            return False
        L, base = self.L, self.L.set_base(self.produced)
        assert n <= L.rows_per_set and payload_bytes <= L.payload_capacity
        w = self.r.write_bytes if hasattr(self.r, "write_bytes") else \
            (lambda off, src, nb, stream=None: self.r.write(off, _as_np(src).view(np.uint8).reshape(-1)[:nb]))
        w(base + L.off_workers, workers, n * 4)
        w(base + L.off_times, times, n * 8)
        w(base + L.off_labels, labels, n * 4)
        w(base + L.off_bitmap, bitmap, n * (L.num_features // 32) * 4)
        w(base + L.off_row_off, row_off, (n + 1) * 4)
        w(base + L.off_payload, payload, payload_bytes)
        hdr = np.zeros(HDR_WORDS, dtype=np.int32)
        hdr[HDR_N], hdr[HDR_PAYLOAD_BYTES], hdr[HDR_SEQ] = n, payload_bytes, self.produced
        self.r.write(base + L.off_hdr, hdr)
        self.produced += 1
        self.r.write(CTL_PRODUCED * 4, np.array([self.produced], dtype=np.int32))     # after the data, same stream
        return True

    def set_flag(self, word: int, value: int = 1) -> None:
        self.r.write(word * 4, np.array([value], dtype=np.int32))

    def worker_state(self):
        """(consumed, worker_done, worker_clock) as last written by the worker app."""
        c = self.r.read(CTL_CONSUMED * 4, 12).view(np.int32)
        self.consumed = int(c[0])
        return int(c[0]), int(c[1]), int(c[2])


@dataclass
class TopicBatch:
    seq: int
    n: int
    payload_bytes: int
    workers: np.ndarray
    times: np.ndarray
    # addresses inside the region (device pointers for a DeviceRegion)
    labels: int
    bitmap: int
    row_off: int
    payload: int


class InputTopicReader:
    """Worker side of the INPUT_DATA topic (the consumer behind ``WorkerSamplingProcessor.process``)."""

    def __init__(self, region, layout: InputTopicLayout):
        self.r, self.L = region, layout
        self.taken = 0
        self.eof = self.stop = False

    def poll(self) -> Optional[TopicBatch]:
        ctl = self.r.read(0, 12).view(np.int32)
        produced, self.eof, self.stop = int(ctl[CTL_PRODUCED]), bool(ctl[CTL_EOF]), bool(ctl[CTL_STOP])
        if produced <= self.taken:
            return None
        L, base = self.L, self.L.set_base(self.taken)
        hdr = self.r.read(base + L.off_hdr, HDR_WORDS * 4).view(np.int32)
        n, nb, seq = int(hdr[HDR_N]), int(hdr[HDR_PAYLOAD_BYTES]), int(hdr[HDR_SEQ])
        if seq != self.taken:
            raise RuntimeError(f"INPUT_DATA slot holds batch {seq}, expected {self.taken}: the topic was overrun")
        workers = self.r.read(base + L.off_workers, n * 4).view(np.int32)
        times = self.r.read(base + L.off_times, n * 8).view(np.float64)
        b = TopicBatch(seq, n, nb, workers, times, self.r.address(base + L.off_labels),
                       self.r.address(base + L.off_bitmap), self.r.address(base + L.off_row_off),
                       self.r.address(base + L.off_payload))
        self.taken += 1
        return b

    @property
    def drained(self) -> bool:
        """The producer is done and every published batch was taken."""
        return self.eof and int(self.r.read(0, 4).view(np.int32)[0]) <= self.taken

    def mark_consumed(self, seq: int, stream=None) -> None:
        self.r.write(CTL_CONSUMED * 4, np.array([seq + 1], dtype=np.int32), stream=stream)

    def set_flag(self, word: int, value: int = 1, stream=None) -> None:
        self.r.write(word * 4, np.array([value], dtype=np.int32), stream=stream)


def _as_np(x) -> np.ndarray:
    return x if isinstance(x, np.ndarray) else x.numpy()


# ------------------------------------------------------------------ rendezvous (the "broker address")
class Rendezvous:
    """TCPStore at the broker address; the server app is the master.  Carries the run description and the IPC
    handles once, at start-up."""

    def __init__(self, address: str, is_server: bool, timeout_s: float = 300.0):
        import datetime
        from torch.distributed import TCPStore
        host, port = address.rsplit(":", 1)
        self.store = TCPStore(host, int(port), world_size=None, is_master=is_server,
                              timeout=datetime.timedelta(seconds=timeout_s), wait_for_workers=False)

    def put(self, key: str, value) -> None:
        self.store.set(key, value if isinstance(value, (bytes, str)) else json.dumps(value))

    def get_bytes(self, key: str) -> bytes:
        return self.store.get(key)

    def get_json(self, key: str):
        return json.loads(self.store.get(key).decode())

    def wait_for(self, key: str, poll_s: float = 0.05, timeout_s: float = 300.0) -> None:
        t0 = time.time()
        while True:
            try:
                if self.store.check([key]):
                    return
            except Exception:
                pass
            if time.time() - t0 > timeout_s:
                raise TimeoutError(f"rendezvous: nobody published {key!r} within {timeout_s:.0f}s")
            time.sleep(poll_s)
