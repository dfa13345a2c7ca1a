"""Property tests (hypothesis) for the host-side protocol pieces: the consistency scheduler under arbitrary
worker interleavings (MessageTracker.java:69-87, ServerProcessor.java:95-134), its C++ twin, the dynamic buffer
window (WorkerSamplingProcessor.java:115-135) and the two wire formats of ingestion."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from psb200.consistency import ConsistencyScheduler, max_gap_bound
from psb200.engine.buffer import DynamicBuffer
from psb200.ops import native

needs_native = pytest.mark.skipif(not native.available(), reason="libpsb200.so not built")


def _run(model, n, picks):
    """Drive a scheduler with an arbitrary schedule: `picks` chooses which READY worker finishes its clock next.
    Returns the largest clock gap ever observed between two workers and the final clocks."""
    sched = ConsistencyScheduler(n, model)
    ready = {w: 0 for w in range(n)}          # worker -> clock it may compute (initial weights at clock 0)
    done = [0] * n                            # gradients pushed per worker
    gap = 0
    for p in picks:
        if not ready:
            break
        w = sorted(ready)[p % len(ready)]
        clock = ready.pop(w)
        assert clock == done[w]
        done[w] += 1
        for rw, rc in sched.on_gradient(w, clock):
            assert rw not in ready, "a worker got two replies for one gradient"
            assert rc == done[rw], "a reply must carry the worker's next clock"
            ready[rw] = rc
        gap = max(gap, max(done) - min(done))
    return gap, done, ready


@settings(max_examples=60, deadline=None)
@given(st.sampled_from([0, 1, 2, 5, -1]), st.integers(2, 6), st.lists(st.integers(0, 1000), min_size=10, max_size=300))
def test_scheduler_respects_its_staleness_bound_and_never_deadlocks(model, n, picks):
    gap, done, ready = _run(model, n, picks)
    bound = max_gap_bound(model)
    if bound is not None:
        assert gap <= bound, (model, gap, bound)
    # liveness: somebody can always proceed (the slowest worker is never blocked)
    assert ready, "no worker may proceed: deadlock"
    slowest = min(range(n), key=lambda w: done[w])
    if model >= 0:
        assert any(done[w] == done[slowest] for w in ready)


@needs_native
@settings(max_examples=40, deadline=None)
@given(st.sampled_from([0, 3, -1]), st.integers(2, 5), st.lists(st.integers(0, 1000), min_size=5, max_size=200))
def test_native_tracker_replies_like_the_python_scheduler(model, n, picks):
    py, cc = ConsistencyScheduler(n, model), native.NativeTracker(n, model)
    ready = {w: 0 for w in range(n)}
    for p in picks:
        if not ready:
            break
        w = sorted(ready)[p % len(ready)]
        clock = ready.pop(w)
        a, b = py.on_gradient(w, clock), cc.on_gradient(w, clock)
        assert sorted(a) == sorted(b)
        ready.update(dict(a))
    assert cc.clocks() == py.tracker.clocks


@settings(max_examples=60, deadline=None)
@given(st.integers(1, 40), st.integers(0, 60), st.floats(0.05, 2.0),
       st.lists(st.floats(0.0, 5000.0, allow_nan=False), min_size=1, max_size=400))
def test_dynamic_buffer_window_invariants(min_size, extra, bc, gaps):
    max_size = min_size + extra
    buf = DynamicBuffer(min_size, max_size, bc)
    t, seen = 0.0, 0
    for g in gaps:
        t += g
        slot = buf.on_event(t)
        seen += 1
        assert 0 <= slot < max_size
        assert min(seen, min_size) <= buf.ring.count <= min(seen, max_size)
        assert min_size <= buf.target <= max_size
        assert buf.ring.head == (slot + 1) % max_size            # the window always ends at the newest row


@needs_native
@settings(max_examples=40, deadline=None)
@given(st.integers(1, 24), st.sampled_from([32, 96, 256]), st.floats(0.0, 1.0), st.integers(1, 40), st.integers(0, 2 ** 31 - 1))
def test_wire_formats_are_lossless(n, F, density, distinct, seed):
    """bitmap + fp16 values and bitmap + palette codes both reproduce arbitrary fp16 rows bit for bit
    (rows with more than 16 distinct values take the raw fallback of the palette format)."""
    rng = np.random.default_rng(seed)
    vocab = rng.standard_normal(distinct).astype(np.float16)
    vocab[vocab == 0] = np.float16(0.5)
    rows = (vocab[rng.integers(0, distinct, (n, F))] * (rng.random((n, F)) < density)).astype(np.float16)
    special = np.array([0x8000, 0x0001, 0x8001, 0x7c00, 0xfc00, 0x7e00, 0xffff], dtype=np.uint16)    # -0, denormals, inf, nan
    hit = rng.random((n, F)) < 0.02
    rows.view(np.uint16)[hit] = special[rng.integers(0, len(special), int(hit.sum()))]
    labels = rng.integers(1, 6, n).astype(np.int32)
    bits = rows.view(np.uint16)
    ref = np.where((bits & 0x7fff) == 0, np.uint16(0), bits)                  # +0 / -0 are absent, all else is kept
    for palette in (False, True):
        pk = native.PackedRows(n, F, pin=False, palette=palette).pack_dense(rows, labels)       # buffers of exact size
        assert np.array_equal(pk.to_dense().view(np.uint16), ref)
        assert pk.labels[:n].tolist() == labels.tolist()
    dense_bytes = n * F * 2
    assert pk.nbytes <= dense_bytes + n * (F // 8 + 8) + 8 + 2 * n          # never much worse than dense
