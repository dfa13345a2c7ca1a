"""Progress watchdog (failure detection).

The reference has no application-level failure detector: a lost gradient trips the
vector-clock assert (``MessageTracker.java:22-35``) and a dead worker blocks the
sequential / bounded models forever (``ServerProcessor.java:111-131``).  Here every
in-kernel wait carries a cycle budget and raises the block's status word
(``csrc/common.cuh: wait_ge_sys``); this host thread polls a progress callback and
the status word and reports who is behind when the clock stops moving.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Optional


class Watchdog:
    def __init__(self, progress: Callable[[], int], status: Callable[[], int], timeout_s: float = 60.0,
                 on_stall: Optional[Callable[[str], None]] = None, poll_s: float = 0.5):
        self.progress, self.status = progress, status
        self.timeout_s, self.poll_s = timeout_s, poll_s
        self.on_stall = on_stall or (lambda msg: print("WATCHDOG: " + msg, flush=True))
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self.tripped: Optional[str] = None

    def start(self) -> "Watchdog":
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        self._thread.join(timeout=2 * self.poll_s + 1)

    def _run(self) -> None:
        last, since = self.progress(), time.time()
        while not self._stop.wait(self.poll_s):
            st = self.status()
            if st != 0:
                self.tripped = f"device wait timed out (status word {st}: 1 = weights never arrived, 2 = clock gate)"
                self.on_stall(self.tripped)
                return
            cur = self.progress()
            if cur != last:
                last, since = cur, time.time()
            elif time.time() - since > self.timeout_s:
                self.tripped = f"no progress for {self.timeout_s:.0f}s at clock {cur}"
                self.on_stall(self.tripped)
                return
