"""Streaming data source: round-robin partitioning and the -p rate limiter.

Reference: ``CsvProducer.java:36-87``: skip header, one row per event, key =
``rowCount % numWorkers`` (round-robin, ``:61``), the first ``numWorkers*128`` rows
are sent unthrottled (``:73``), afterwards the thread sleeps 1000 ms every
``1000/p`` rows (``:75-82``) => ``1000/p`` rows/s *in total*; prints
``CSV Producer sent all records!`` at EOF (``:85``).

Quirk Q7: the integer division makes ``p > 1000`` a modulo by zero there; here
``p > 1000`` means "sleep p ms per row" and ``p == 0`` means unthrottled.

``PacedSchedule`` is the pure timing rule on a virtual clock (what the C++
producer thread in ``csrc/host_producer.cpp`` executes against the wall clock).
"""
from __future__ import annotations

from typing import Iterator, Tuple

EOF_LINE = "CSV Producer sent all records!"


class PacedSchedule:
    def __init__(self, num_workers: int, time_per_event_ms: int, warmup_rows_per_worker: int = 128):
        self.num_workers = num_workers
        self.p = int(time_per_event_ms)
        self.warmup = num_workers * warmup_rows_per_worker
        self.row_count = 0
        self.now_ms = 0.0

    def next(self) -> Tuple[int, float]:
        """Returns (partition, virtual send time in ms) for the next row and advances."""
        part = self.row_count % self.num_workers
        t = self.now_ms
        self.row_count += 1
        if self.p > 0 and self.row_count >= self.warmup:
            if self.p <= 1000:
                per_sec = 1000 // self.p
                if self.row_count % per_sec == 0:
                    self.now_ms += 1000.0
            else:
                self.now_ms += float(self.p)
        return part, t

    def sleep_after(self, row_count_after: int) -> float:
        """Sleep (ms) the producer takes after having sent `row_count_after` rows."""
        if self.p <= 0 or row_count_after < self.warmup:
            return 0.0
        if self.p <= 1000:
            return 1000.0 if row_count_after % (1000 // self.p) == 0 else 0.0
        return float(self.p)


def schedule(num_rows: int, num_workers: int, time_per_event_ms: int) -> Iterator[Tuple[int, int, float]]:
    s = PacedSchedule(num_workers, time_per_event_ms)
    for i in range(num_rows):
        part, t = s.next()
        yield i, part, t
