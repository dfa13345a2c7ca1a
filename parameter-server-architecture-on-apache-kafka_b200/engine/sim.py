"""Event-driven parameter-server engine (producer + workers + server in one process).

This is the closed loop of the reference — ``WorkerTrainingProcessor.process``
(``WorkerTrainingProcessor.java:63-98``) <-> ``ServerProcessor.process``
(``ServerProcessor.java:143-183``) fed by ``CsvProducer`` / ``WorkerSamplingProcessor``
— with the three Kafka topics replaced by an in-process event queue.  It is
exact in semantics (same scheduler, same buffer policy, same log rows) and runs
either on a *virtual* clock (deterministic; worker iteration time is a model
parameter, which is what makes bounded-delay / eventual runs reproducible and
lets tests inject stragglers) or against the wall clock.

The math is delegated to a backend: ``NumpyBackend`` (CPU oracle) or
``psb200.engine.cuda_backend.CudaBackend`` (sm_100a kernels).  The throughput
path on many GPUs is ``psb200.engine.device_engine`` instead; this engine is what
the two Runner CLIs drive for reference-rate streaming runs and what every
protocol test uses.
"""
from __future__ import annotations

import heapq
import itertools
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np

from ..config import PSConfig
from ..consistency import ConsistencyScheduler
from ..data.csr import CSRBatch
from ..models.logreg import LogRegLayout
from ..oracle.lr_task import LogisticRegressionTask
from ..utils.logs import LogSink, server_row, worker_row
from ..utils.metrics import Metrics
from .buffer import DynamicBuffer
from .producer import EOF_LINE, PacedSchedule

VIRTUAL_EPOCH_MS = 1_584_302_400_000      # timestamps of virtual-clock runs start here


# --------------------------------------------------------------------------- backends
class NumpyBackend:
    """CPU oracle backend: one ``LogisticRegressionTask`` per worker + one for the server."""

    name = "numpy"

    def __init__(self, cfg: PSConfig, test: Optional[CSRBatch]):
        self.cfg = cfg
        self.layout = LogRegLayout(cfg.num_features, cfg.num_classes)
        mk = lambda: LogisticRegressionTask(cfg.num_features, cfg.num_classes, test, cfg.local_solver,
                                            cfg.max_iter, cfg.sgd_lr, cfg.standardize, cfg.center)
        self.tasks = [mk() for _ in range(cfg.num_workers)]
        self.server_task = mk()
        self.rows: List[Dict[int, Tuple[np.ndarray, np.ndarray, int]]] = [dict() for _ in range(cfg.num_workers)]
        self.weights = self.layout.zeros() if cfg.init == "zero" else self.layout.random(cfg.seed)

    def store_row(self, worker: int, slot: int, idx: np.ndarray, val: np.ndarray, label: int) -> None:
        self.rows[worker][slot] = (idx, val, label)

    def _buffer(self, worker: int, slots: List[int]) -> CSRBatch:
        rows = [self.rows[worker][s] for s in slots]
        lens = [len(r[0]) for r in rows]
        indptr = np.zeros(len(rows) + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        return CSRBatch(indptr, np.concatenate([r[0] for r in rows]).astype(np.int32),
                        np.concatenate([r[1] for r in rows]).astype(np.float32),
                        np.asarray([r[2] for r in rows], dtype=np.int32), self.cfg.num_features)

    def snapshot(self) -> np.ndarray:
        return self.weights.copy()

    def worker_step(self, worker: int, weights: np.ndarray, slots: List[int], want_metrics: bool):
        t = self.tasks[worker]
        t.set_weights(weights)
        delta = t.calculate_gradients(self._buffer(worker, slots), eval_test=want_metrics)
        return delta, t.loss, (t.metrics if want_metrics else None)

    def apply(self, delta: np.ndarray, lr: float) -> None:
        if self.cfg.compat_q1:                # key 6149 (last intercept) is never updated (Q1)
            delta = delta.copy()
            delta[self.layout.F, self.layout.R - 1] = 0.0
        self.weights += np.float32(lr) * delta

    def server_metrics(self) -> Optional[Metrics]:
        self.server_task.set_weights(self.weights)
        if self.server_task._Xtest is None:
            return None
        return self.server_task.calculate_test_metrics()


# --------------------------------------------------------------------------- engine
@dataclass(order=True)
class _Event:
    time: float
    seq: int
    kind: str = field(compare=False)
    payload: tuple = field(compare=False, default=())


class PSSimulation:
    def __init__(self, cfg: PSConfig, train: CSRBatch, test: Optional[CSRBatch] = None, backend=None,
                 server_log: Optional[LogSink] = None, worker_log: Optional[LogSink] = None,
                 iter_time_ms: Optional[Callable[[int, int], float]] = None, virtual: bool = True,
                 server_start_ms: float = 20000.0, worker_metrics: bool = True, link=None):
        self.cfg = cfg
        self.train, self.test = train, test
        self.backend = backend if backend is not None else NumpyBackend(cfg, test)
        self.server_log = server_log or LogSink(keep=True, stream=_Null())
        self.worker_log = worker_log or LogSink(keep=True, stream=_Null())
        self.sched = ConsistencyScheduler(cfg.num_workers, cfg.consistency_model)
        self.buffers = [DynamicBuffer(cfg.min_buffer_size, cfg.max_buffer_size, cfg.buffer_size_coefficient)
                        for _ in range(cfg.num_workers)]
        self.iter_time_ms = iter_time_ms or (lambda worker, clock: 2700.0)
        self.virtual = virtual
        self.server_start_ms = server_start_ms      # ServerAppRunner.java:95 sleeps 20 s before the loop starts
        self.worker_metrics = worker_metrics
        self.link = link                    # psb200.parallel.link.Link to a remote worker app, or None (inline workers)
        self.outstanding = 0                # remote worker steps in flight
        self._delays = _parse_delays(cfg.inject_delay)
        self._q: List[_Event] = []
        self._seq = itertools.count()
        self.now = 0.0
        self.server_clock = 0
        self.started: Dict[int, int] = {}          # worker -> clock in progress
        self.max_gap = 0
        self.rows_sent = 0
        self.producer_done = False
        self._stop_clock = 0

    # -- clock / queue ------------------------------------------------------
    def _push(self, t: float, kind: str, *payload) -> None:
        heapq.heappush(self._q, _Event(t, next(self._seq), kind, payload))

    def _stamp(self) -> int:
        return int(VIRTUAL_EPOCH_MS + self.now) if self.virtual else int(time.time() * 1000)

    # -- run ------------------------------------------------------------------
    def run(self, max_clock: Optional[int] = None, max_virtual_ms: Optional[float] = None) -> None:
        cfg = self.cfg
        max_clock = cfg.max_clock if max_clock is None else max_clock
        sched = PacedSchedule(cfg.num_workers, cfg.producer_time_per_event)
        for i in range(len(self.train)):
            part, t = sched.next()
            self._push(t, "row", i, part)
        self._push(sched.now_ms, "eof")
        self._push(self.server_start_ms + 1000.0, "start")     # 1 s punctuator, ServerProcessor.java:61
        t0 = time.time()
        while True:
            if self.link is not None:
                self._poll_link()
            if not self._q:
                if self.link is not None and self.link.outbox:
                    self._poll_link(force=True)
                    continue
                break
            ev = self._q[0]
            if max_virtual_ms is not None and ev.time > max_virtual_ms:
                break
            if not self.virtual:
                wait = ev.time / 1000.0 - (time.time() - t0)
                if wait > 0:
                    time.sleep(min(wait, 0.001 if self.link is not None else wait))
                    continue
            heapq.heappop(self._q)
            self.now = ev.time if self.virtual else max(ev.time, (time.time() - t0) * 1000.0)
            getattr(self, "_on_" + ev.kind)(*ev.payload)
            if max_clock is not None and max_clock >= 0 and self.server_clock >= max_clock:
                break
            if (max_clock is None or max_clock < 0) and self.producer_done and self.server_clock >= self._stop_clock:
                break           # the reference never stops; we drain one more clock after EOF
        if self.link is not None:
            from ..parallel.link import STOP
            self.link.send(STOP)
            self._poll_link(force=True)

    def _poll_link(self, force: bool = False) -> None:
        """Lock-step exchange with the worker app: ship queued ROW / WEIGHTS messages, collect gradients."""
        from ..parallel.link import GRADIENT
        if not self.link.outbox or not (force or self.link.pending_weights() or len(self.link.outbox) >= 256):
            return
        for kind, worker, clock, ints, floats, aux in self.link.exchange():
            if kind != GRADIENT:
                raise RuntimeError(f"server app received unexpected message kind {kind}")
            n = self.backend.layout.device_size
            delta = floats[:n].reshape(self.backend.layout.device_rows, self.backend.layout.CP).copy()
            loss, f1, acc = (float(v) for v in floats[n:n + 3])
            self.outstanding -= 1
            self._push(self.now, "grad", worker, clock, delta, loss, f1, acc, int(aux))

    # -- handlers -------------------------------------------------------------
    def _on_row(self, i: int, part: int) -> None:
        idx, val, lab = self.train.row(i)
        self.rows_sent += 1
        if self.link is not None:           # the remote sampling processor owns the buffer
            from ..parallel.link import ROW
            self.link.send(ROW, part, 0, idx, val, aux=lab)
            return
        slot = self.buffers[part].on_event(self.now)
        self.backend.store_row(part, slot, idx, val, lab)

    def _on_eof(self) -> None:
        self.producer_done = True
        self._stop_clock = self.server_clock + 1
        self.server_log.println(EOF_LINE)                # lands in logs-server.csv with -l (CsvProducer.java:85)

    def _on_start(self) -> None:
        snap = self.backend.snapshot()
        for w in range(self.cfg.num_workers):
            self._push(self.now, "wstart", w, 0, snap)

    def _on_wstart(self, worker: int, clock: int, weights: np.ndarray) -> None:
        if self.link is not None:
            from ..parallel.link import WEIGHTS
            self.started[worker] = clock
            self.max_gap = max(self.max_gap, max(self.started.values()) - min(self.started.values()))
            self.link.send(WEIGHTS, worker, clock, None, weights)
            self.outstanding += 1
            return
        buf = self.buffers[worker]
        slots = buf.ring.active_slots()
        if not slots:
            raise RuntimeError(f"There is no data for partition {worker}")    # WorkerTrainingProcessor.java:131-133
        self.started[worker] = clock
        self.max_gap = max(self.max_gap, max(self.started.values()) - min(self.started.values()))
        delta, loss, m = self.backend.worker_step(worker, weights, slots, self.worker_metrics)
        f1, acc = (m.f1, m.accuracy) if m is not None else (-1.0, -1.0)
        dur = self.iter_time_ms(worker, clock) + self._delays.get(worker, 0.0)
        self._push(self.now + dur, "grad", worker, clock, delta, loss, f1, acc, buf.ring.total)

    def _on_grad(self, worker: int, clock: int, delta, loss, f1, acc, seen) -> None:
        self.worker_log.println(worker_row(self._stamp(), worker, clock, loss, f1, acc, seen))
        self.backend.apply(delta, self.cfg.server_learning_rate)
        replies = self.sched.on_gradient(worker, clock)
        if worker == 0:                                   # Q6: one server row per clock of worker 0
            self.server_clock = clock + 1
            if clock % max(1, self.cfg.eval_every) == 0:
                m = self.backend.server_metrics()
                if m is not None:
                    self.server_log.println(server_row(self._stamp(), clock, m.f1, m.accuracy))
        if replies:
            snap = self.backend.snapshot()
            for q, vc in replies:
                self._push(self.now, "wstart", q, vc, snap)


class _Null:
    def write(self, s):
        return len(s)

    def flush(self):
        pass


def _parse_delays(spec: str) -> Dict[int, float]:
    out: Dict[int, float] = {}
    for item in filter(None, (spec or "").split(",")):
        r, ms = item.split(":")
        out[int(r)] = float(ms)
    return out
