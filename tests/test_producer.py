"""Producer pacing and round-robin partitioning (CsvProducer.java:61-83)."""
from psb200.engine.producer import PacedSchedule, schedule


def test_round_robin_and_warmup_then_rate():
    rows = list(schedule(4 * 128 + 50, 4, 200))
    assert [p for _, p, _ in rows[:8]] == [0, 1, 2, 3, 0, 1, 2, 3]
    assert all(t == 0.0 for _, _, t in rows[:4 * 128])                 # first numWorkers*128 rows unthrottled
    tail = [t for _, _, t in rows[4 * 128:]]
    # 1000/200 = 5 rows per second in total afterwards
    assert tail[0] == 0.0 or tail[0] == 1000.0
    assert tail[-1] - tail[0] == 1000.0 * ((len(tail) - 1) // 5) or tail[-1] - tail[0] == 1000.0 * (len(tail) // 5)


def test_unthrottled_and_large_p():
    assert all(t == 0.0 for _, _, t in schedule(2000, 4, 0))
    s = PacedSchedule(1, 1500)
    times = [s.next()[1] for _ in range(140)]
    assert times[127] == 0.0 and times[130] - times[129] == 1500.0     # p > 1000: sleep p ms per row (Q7)


def test_rate_is_total_not_per_worker():
    n = 2 * 128 + 1000
    rows = list(schedule(n, 2, 100))                                   # 10 rows/s total
    span_s = (rows[-1][2] - rows[2 * 128][2]) / 1000.0
    assert abs(1000 / span_s - 10.0) < 0.2
