"""Vector-clock bookkeeping and the three consistency schedulers.

Behavioural contract (reference): ``MessageTracker.java:10-88`` keeps, per worker,
the count of gradients received (= next clock to hand out) and a flag saying
whether the weights for that clock were already sent; ``ServerProcessor.java:95-134``
turns a received gradient into the set of workers to answer:

* eventual (-1): answer the sender at once with clock+1;
* sequential (0): when every worker has delivered clock v, answer all with v+1;
* bounded delay (k>0): answer every un-answered worker q whose clock is at most
  k ahead of the slowest one (``min clock >= clock[q] - k``).

This module is the host-side, pure-Python statement of that contract.  The same
predicate is evaluated inside the kernels on clock words in peer memory
(``csrc/comm.cuh``: ``comm_pull`` / ``wait_ge_sys``) and in the C++ runtime
(``csrc/host_runtime.cpp``); tests check all three against each other.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

from ..config import CONSISTENCY_EVENTUAL, CONSISTENCY_SEQUENTIAL


class ProtocolViolation(ValueError):
    """A gradient or a reply carried an unexpected clock (lost / duplicated / reordered)."""


@dataclass
class WorkerStatus:
    clock: int = 0          # gradients received so far == clock the worker works on next
    sent: bool = True       # weights for `clock` already handed out (initial broadcast counts)


class ClockTracker:
    def __init__(self, num_workers: int):
        if num_workers < 1:
            raise ValueError("num_workers must be >= 1")
        self.num_workers = num_workers
        self.status: List[WorkerStatus] = [WorkerStatus() for _ in range(num_workers)]

    # -- mutation ---------------------------------------------------------
    def received(self, worker: int, clock: int) -> None:
        st = self.status[worker]
        if st.clock != clock:
            raise ProtocolViolation(f"Expected value {st.clock}, actual value {clock}")
        st.clock += 1
        st.sent = False

    def sent(self, worker: int, clock: int) -> None:
        st = self.status[worker]
        if st.clock != clock:
            raise ProtocolViolation(f"Expected value {st.clock}, actual value {clock}")
        st.sent = True

    def sent_all(self, clock: int) -> None:
        for w in range(self.num_workers):
            self.sent(w, clock)

    # -- queries ----------------------------------------------------------
    @property
    def clocks(self) -> List[int]:
        return [s.clock for s in self.status]

    def min_clock(self) -> int:
        return min(s.clock for s in self.status)

    def has_received_all(self, clock: int) -> bool:
        return self.min_clock() >= clock + 1

    def sendable(self, max_delay: int) -> List[Tuple[int, int]]:
        lo = self.min_clock()
        return [(w, s.clock) for w, s in enumerate(self.status)
                if not s.sent and lo >= s.clock - max_delay]


class ConsistencyScheduler:
    """received gradient -> [(worker, clock)] replies, and marks them sent."""

    def __init__(self, num_workers: int, consistency_model: int):
        if consistency_model < -1:
            raise ValueError("consistency_model must be -1 (eventual), 0 (sequential) or k>0 (bounded delay)")
        self.model = consistency_model
        self.tracker = ClockTracker(num_workers)

    @property
    def name(self) -> str:
        if self.model == CONSISTENCY_EVENTUAL:
            return "eventual"
        if self.model == CONSISTENCY_SEQUENTIAL:
            return "sequential"
        return f"bounded_delay_{self.model}"

    def on_gradient(self, worker: int, clock: int) -> List[Tuple[int, int]]:
        t = self.tracker
        t.received(worker, clock)
        if self.model == CONSISTENCY_EVENTUAL:
            replies = [(worker, clock + 1)]
        elif self.model == CONSISTENCY_SEQUENTIAL:
            replies = ([(w, clock + 1) for w in range(t.num_workers)]
                       if t.has_received_all(clock) else [])
        else:
            replies = t.sendable(self.model)
        for w, c in replies:
            t.sent(w, c)
        return replies

    def may_start(self, worker: int, clock: int) -> bool:
        """The same rule as a predicate: may `worker` begin iteration `clock`?

        This is the form the kernels spin on: ``min_q clock[q] >= clock - k``
        with k=0 (sequential) or k=inf (eventual).  ``clock[q]`` counts the
        gradients of q already applied.
        """
        if clock == 0 or self.model == CONSISTENCY_EVENTUAL:
            return True
        return self.tracker.min_clock() >= clock - self.model


def max_gap_bound(consistency_model: int):
    """Largest in-progress clock difference the model permits (None: unbounded)."""
    if consistency_model == CONSISTENCY_EVENTUAL:
        return None
    return consistency_model + 1 if consistency_model > 0 else 1
