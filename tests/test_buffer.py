"""Dynamic buffer: sizing rule and O(1) ring vs the literal slot-map policy (WorkerSamplingProcessor.java:50-135)."""
import random

from psb200.engine.buffer import (DynamicBuffer, InterArrivalStats, ReferenceStore, RingWindow,
                                  target_buffer_size)


def test_target_defaults_and_clamps():
    assert target_buffer_size(1000.0, 0.3, 128, 1024) == 128          # default delta: 60 ev/min * 0.3 = 18 -> min
    assert target_buffer_size(100.0, 0.3, 128, 1024) == 180           # 600 ev/min * 0.3
    assert target_buffer_size(1.0, 0.3, 128, 1024) == 1024            # clamps to max
    assert target_buffer_size(0.0, 0.3, 128, 1024) == 1024            # Q8 fixed: infinite rate -> max
    assert target_buffer_size(0.0, 0.3, 128, 1024, compat_q8=True) == 128
    assert target_buffer_size(50.0, 0.3, 128, 1024) == 360


def test_interarrival_window_of_500():
    st = InterArrivalStats()
    assert st.mean_delta_ms() == 1000.0
    t = 0.0
    st.observe(t)
    assert st.mean_delta_ms() == 1000.0                                 # first event only sets the reference time
    for i in range(700):
        t += 10.0 if i < 200 else 20.0
        st.observe(t)
    assert len(st.deltas) == 500 and abs(st.mean_delta_ms() - 20.0) < 1e-9


def test_ring_equals_reference_policy_under_changing_targets():
    rng = random.Random(1)
    cap = 64
    ring, ref = RingWindow(cap), ReferenceStore(cap)
    target = 8
    for step in range(2000):
        if step % 37 == 0:
            target = rng.randint(1, cap)
        prev = ring.count
        ring.push(target)
        ref.push(target)
        assert ring.count == min(prev + 1, target)
        lo = ring.total - ring.count + 1
        assert ref.active_ids() == list(range(lo, ring.total + 1))     # the most recent `count` insertions
        assert len(set(ring.active_slots())) == ring.count
        assert all(ring.is_active(s) for s in ring.active_slots())


def test_dynamic_buffer_tracks_rate():
    b = DynamicBuffer(128, 1024, 0.3)
    t = 0.0
    for _ in range(2000):
        t += 50.0                       # 1200 events/min -> target 360
        b.on_event(t)
    assert b.target == 360 and b.ring.count == 360 and b.ring.total == 2000
    for _ in range(600):
        t += 1000.0                     # 60 events/min -> shrinks to min as the window forgets
        b.on_event(t)
    assert b.target == 128 and b.ring.count == 128
