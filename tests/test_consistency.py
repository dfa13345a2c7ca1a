"""Consistency tracker and scheduler vs the reference rules (MessageTracker.java, ServerProcessor.java:95-134)."""
import random

import pytest

from psb200.consistency import ClockTracker, ConsistencyScheduler, ProtocolViolation, max_gap_bound


def test_initial_state_and_asserts():
    t = ClockTracker(3)
    assert t.clocks == [0, 0, 0] and all(s.sent for s in t.status)
    t.received(1, 0)
    assert t.clocks == [0, 1, 0] and not t.status[1].sent
    with pytest.raises(ProtocolViolation):
        t.received(1, 0)            # duplicate
    with pytest.raises(ProtocolViolation):
        t.received(0, 1)            # skipped a clock
    with pytest.raises(ProtocolViolation):
        t.sent(1, 0)                # reply must carry the next clock
    t.sent(1, 1)
    assert t.status[1].sent


def test_has_received_all():
    t = ClockTracker(2)
    assert not t.has_received_all(0)
    t.received(0, 0)
    assert not t.has_received_all(0)
    t.received(1, 0)
    assert t.has_received_all(0) and not t.has_received_all(1)


def test_sequential_replies_only_when_all_arrived():
    s = ConsistencyScheduler(3, 0)
    assert s.on_gradient(0, 0) == []
    assert s.on_gradient(2, 0) == []
    assert sorted(s.on_gradient(1, 0)) == [(0, 1), (1, 1), (2, 1)]


def test_eventual_replies_to_sender_immediately():
    s = ConsistencyScheduler(3, -1)
    for c in range(5):
        assert s.on_gradient(1, c) == [(1, c + 1)]
    assert s.tracker.clocks == [0, 5, 0]


def test_bounded_delay_blocks_at_k_and_releases_on_slow_push():
    k = 2
    s = ConsistencyScheduler(2, k)
    # worker 0 runs ahead: clocks 0..k are answered, k+1-th gradient is not
    for c in range(k):
        assert s.on_gradient(0, c) == [(0, c + 1)]
    assert s.on_gradient(0, k) == []               # clock[0]=k+1 > min(0)+k  -> withheld
    # the slow worker's push releases the fast one (and itself)
    assert sorted(s.on_gradient(1, 0)) == [(0, k + 1), (1, 1)]


def _simulate(num_workers, model, iters, speeds, seed=0):
    """Event simulation with per-worker iteration times; returns max in-progress clock gap."""
    import heapq
    s = ConsistencyScheduler(num_workers, model)
    q = [(speeds[w], w, 0) for w in range(num_workers)]
    heapq.heapify(q)
    progress = {w: 0 for w in range(num_workers)}
    gap = 0
    done = 0
    while q and done < iters:
        t, w, c = heapq.heappop(q)
        done += 1
        for (r, vc) in s.on_gradient(w, c):
            progress[r] = vc
            heapq.heappush(q, (t + speeds[r], r, vc))
        gap = max(gap, max(progress.values()) - min(progress.values()))
    return gap, done


@pytest.mark.parametrize("model", [0, 1, 3, 10])
def test_gap_bounds_hold(model):
    gap, done = _simulate(3, model, 600, [1.0, 1.7, 9.0])
    assert done == 600                                 # no deadlock
    assert gap <= max_gap_bound(model)
    if model > 0:
        assert gap >= model                            # the straggler makes the bound (nearly) tight


def test_eventual_gap_unbounded_and_no_deadlock():
    gap, done = _simulate(3, -1, 600, [1.0, 1.7, 9.0])
    assert done == 600 and gap > 11 and max_gap_bound(-1) is None


def test_may_start_predicate_matches_replies():
    rng = random.Random(3)
    for model in (0, 2, -1):
        s = ConsistencyScheduler(4, model)
        waiting = {w: 0 for w in range(4)}           # worker -> clock it may run (already replied)
        blocked = {}
        for _ in range(400):
            w = rng.choice(list(waiting))
            c = waiting.pop(w)
            blocked[w] = c + 1
            for r, vc in s.on_gradient(w, c):
                assert blocked.pop(r) == vc
                waiting[r] = vc
            for r, vc in blocked.items():            # withheld replies are exactly the ones the predicate denies
                assert not s.may_start(r, vc)
            for r, vc in waiting.items():
                assert s.may_start(r, vc)
            if not waiting:
                break
