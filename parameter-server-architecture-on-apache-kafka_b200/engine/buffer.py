"""Dynamic input buffer: sizing rule and eviction policy.

Reference: ``WorkerSamplingProcessor.java:50-135``.  Per event the reference
(1) records the inter-arrival time (window of the last 500 deltas, ``:124-135``),
(2) computes ``target = clamp(round(bc * events_per_minute), min, max)`` with
``events_per_minute = 60000 / mean_delta_ms`` and a default delta of 1000 ms
(``:115-122``), (3) scans its key range of the state store and grows / replaces
the oldest / shrinks (``:81-107``) — O(buffer) JSON decodes per row.

Net effect of (3): after each event the buffer holds the most recent
``min(previous_size + 1, target)`` rows.  ``RingWindow`` implements exactly that
in O(1) on a circular slot array (the device ring, ``csrc/ps_ring.cu``);
``ReferenceStore`` restates the slot-map algorithm literally and is the oracle
the tests compare the ring against.

Quirk Q8: an infinite event rate makes the reference's ``(int) Math.round(+Inf)``
wrap to -1 and clamp to *min*; here it clamps to *max* unless ``compat_q8``.
"""
from __future__ import annotations

import math
from collections import deque
from typing import Dict, List, Optional

from ..config import PROCESSING_INTERVAL_SIZE


class InterArrivalStats:
    def __init__(self, window: int = PROCESSING_INTERVAL_SIZE):
        self.window = window
        self.deltas: deque = deque()
        self.total = 0.0
        self.last: Optional[float] = None

    def observe(self, now_ms: float) -> None:
        if self.last is None:
            self.last = now_ms
            return
        d = now_ms - self.last
        self.last = now_ms
        self.deltas.append(d)
        self.total += d
        if len(self.deltas) > self.window:
            self.total -= self.deltas.popleft()

    def mean_delta_ms(self) -> float:
        return self.total / len(self.deltas) if self.deltas else 1000.0


def target_buffer_size(mean_delta_ms: float, bc: float, min_size: int, max_size: int,
                       compat_q8: bool = False) -> int:
    if mean_delta_ms <= 0:
        calc = -1 if compat_q8 else max_size
    else:
        epm = 60000.0 / mean_delta_ms
        v = bc * epm
        calc = int(math.floor(v + 0.5)) if math.isfinite(v) else (-1 if compat_q8 else max_size)   # Math.round
    return max(min_size, min(max_size, calc))


class RingWindow:
    """O(1) sliding window over a circular slot array of ``capacity`` slots.

    ``push`` returns the slot to write.  Active rows are the last ``count``
    insertions, i.e. slots ``(head - count .. head - 1) mod capacity``.
    """

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.head = 0           # next slot to write
        self.count = 0
        self.total = 0          # rows ever inserted == largest insertion id

    def push(self, target: int) -> int:
        if not (1 <= target <= self.capacity):
            raise ValueError("target out of range")
        slot = self.head
        self.head = (self.head + 1) % self.capacity
        self.count = min(self.count + 1, target)
        self.total += 1
        return slot

    def active_slots(self) -> List[int]:
        return [(self.head - self.count + i) % self.capacity for i in range(self.count)]

    def is_active(self, slot: int) -> bool:
        return ((self.head - 1 - slot) % self.capacity) < self.count


class ReferenceStore:
    """Literal restatement of the slot-map policy (one partition's key range)."""

    def __init__(self, capacity: int):
        self.capacity = capacity
        self.slots: Dict[int, int] = {}          # key -> insertion id

    def push(self, target: int) -> int:
        index = list(self.slots.items())
        largest = max((v for _, v in index), default=0)
        if len(index) < target:
            key = min(k for k in range(self.capacity) if k not in self.slots)
        elif len(index) == target:
            key = min(index, key=lambda kv: kv[1])[0]
        else:
            index.sort(key=lambda kv: kv[1])
            n_remove = len(index) - target
            for i in range(n_remove):
                del self.slots[index[i][0]]
            key = index[n_remove][0]
        self.slots[key] = largest + 1
        return key

    def active_ids(self) -> List[int]:
        return sorted(self.slots.values())


class DynamicBuffer:
    """Sizing rule + ring window for one worker (host mirror of the device ring)."""

    def __init__(self, min_size: int, max_size: int, bc: float, compat_q8: bool = False):
        self.min_size, self.max_size, self.bc, self.compat_q8 = min_size, max_size, bc, compat_q8
        self.stats = InterArrivalStats()
        self.ring = RingWindow(max_size)
        self.target = min_size

    def on_event(self, now_ms: float) -> int:
        self.stats.observe(now_ms)
        self.target = target_buffer_size(self.stats.mean_delta_ms(), self.bc, self.min_size,
                                         self.max_size, self.compat_q8)
        return self.ring.push(self.target)
