"""Per-worker consistency inside the fused engine (``PSB_COMM_WORKER``): every worker CTA owns a clock word, applies
its own delta on arrival and is gated on ``min over all workers' clocks >= v - k`` — the granularity of the
reference's tracker (``MessageTracker.java:69-87``, ``ServerProcessor.java:95-134``).  The gates here really block:
the workers of a launch (and the launches of several virtual ranks, on separate streams) are on the GPU together."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.config import PSConfig
from psb200.data import ReviewGenerator

pytestmark = pytest.mark.gpu

F, CAP, N = 256, 128, 96
SLOW = 3_000_000            # SM cycles the straggler sleeps before each of its clocks (~1.5 ms; a clock takes ~0.1 ms)


def _fill(engines, seed=0):
    gen = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0, seed=seed)
    for e in engines:
        b = gen.train_split(N * e.W, seed=seed + 10 + e.rank)
        rows = torch.from_numpy(b.to_dense(np.float16)).view(e.W, N, F)
        labs = torch.from_numpy(b.labels.astype(np.int32)).view(e.W, N)
        e.fill(rows.cuda(), labs.cuda())


def _engine(model, W, impl, **kw):
    from psb200.engine.device_engine import DeviceEngine
    cfg = PSConfig(num_workers=W, num_features=F, consistency_model=model, max_buffer_size=CAP, min_buffer_size=16)
    return DeviceEngine(cfg, W, impl=impl, async_granularity="worker", spin_limit_cycles=4_000_000_000, **kw)


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("k", [1, 3])
def test_bounded_delay_holds_between_the_workers_of_one_gpu(impl, k):
    clocks, W = 12, 4
    e = _engine(k, W, impl)
    try:
        _fill([e])
        e.inject_delay(0, SLOW)                               # worker 0 is the straggler
        e.step(num_clocks=clocks)
        h = e.read_hist()
        assert e.status() == 0
        assert (e.worker_clocks() == clocks).all()            # everybody finished every clock
        assert np.isfinite(h["loss"]).all()
        assert (h["clock"] == np.arange(clocks)[:, None]).all()
        assert h["lag"].max() == k                            # the fast workers ran into the gate: exactly k ahead ...
        assert h["lag"].min() >= 0 and h["lag"][:, 0].max() <= 1   # ... of the straggler, which never waits for anyone
        assert h["lead"][:, 0].max() >= k                     # seen from the straggler the others are k (+1 finishing) ahead
        assert h["lead"].max() <= k + 1
    finally:
        e.close()


@pytest.mark.parametrize("impl", [0, 1])
def test_eventual_workers_of_one_gpu_run_away_from_a_straggler(impl):
    clocks, W = 12, 4
    e = _engine(-1, W, impl)
    try:
        _fill([e])
        e.inject_delay(0, SLOW)
        e.step(num_clocks=clocks)
        h = e.read_hist()
        assert e.status() == 0 and (e.worker_clocks() == clocks).all() and np.isfinite(h["loss"]).all()
        assert h["lag"][:, 1:].max() >= clocks - 3            # nobody waits: the fast ones finish while worker 0 has barely begun
        assert h["loss"][-1].mean() < h["loss"][0].mean()
    finally:
        e.close()


@pytest.mark.parametrize("impl", [0, 1])
def test_one_worker_per_worker_mode_equals_the_sequential_engine(impl):
    """With a single worker there is no concurrency: RED lr*delta into the shard must give the lockstep result."""
    from psb200.engine.device_engine import DeviceEngine
    cfg0 = PSConfig(num_workers=1, num_features=F, consistency_model=0, max_buffer_size=CAP, min_buffer_size=16)
    a = DeviceEngine(cfg0, 1, impl=impl)
    b = _engine(-1, 1, impl)
    try:
        _fill([a]), _fill([b])
        la = []
        for _ in range(3):
            a.step()
            la.append(float(a.read_loss()[0]))
        b.step(num_clocks=3)
        hb = b.read_hist()
        assert np.allclose(hb["loss"][:, 0], la, rtol=2e-4, atol=2e-5)
        wa, wb = a.weights().cpu().numpy(), b.weights().cpu().numpy()
        assert np.abs(wa - wb).max() < 2e-4 * max(1e-2, np.abs(wa).max())
    finally:
        a.close(), b.close()


def test_two_virtual_ranks_share_one_replicated_clock_table():
    """Two ranks' launches on two streams of one GPU, eventual consistency (nobody waits, so the two launches need
    not overlap; launches that wait on one another must not share a GPU — the bounded-delay gate across ranks is
    tested on real GPUs, tests/test_gpu_multi.py): every worker publishes its clocks into BOTH blocks."""
    from psb200.engine.device_engine import VirtualCluster
    clocks, wpr = 10, 2
    cfg = PSConfig(num_workers=2 * wpr, num_features=F, consistency_model=-1, max_buffer_size=CAP, min_buffer_size=16)
    vc = VirtualCluster(cfg, 2, wpr, 1, async_granularity="worker", spin_limit_cycles=4_000_000_000)
    try:
        _fill(vc.engines)
        vc.step_concurrent(num_clocks=clocks)
        hs = [e.read_hist() for e in vc.engines]
        assert all(e.status() == 0 for e in vc.engines)
        for e in vc.engines:
            assert (e.worker_clocks() == clocks).all()        # both copies of the table are complete
        assert all(np.isfinite(h["loss"]).all() for h in hs)
        w = [e.weights().cpu().numpy() for e in vc.engines]   # each rank reads all shards from their owners
        assert np.abs(w[0] - w[1]).max() < 1e-6 and np.abs(w[0]).max() > 0
    finally:
        vc.close()


def test_groups_of_one_gpu_advance_on_their_own_streams():
    """groups=2: two launches per step on two streams; a delayed group does not hold back the other one's launches."""
    clocks, W = 6, 4
    e = _engine(-1, W, 1, groups=2)
    try:
        _fill([e])
        e.inject_delay(2, SLOW)                               # a worker of group 1
        for _ in range(2):
            e.step_group(0, clocks)                           # group 0 runs 12 clocks ...
        e.step_group(1, clocks)                               # ... while group 1 runs 6
        torch.cuda.synchronize()
        c = e.worker_clocks()
        assert c[:2].tolist() == [12, 12] and c[2:].tolist() == [6, 6] and e.status() == 0
        assert np.isfinite(e.read_hist(0)["loss"]).all() and np.isfinite(e.read_hist(1)["loss"]).all()
    finally:
        e.close()


def test_gated_launch_that_cannot_be_resident_is_refused():
    from psb200.ops import native
    e = _engine(2, 1, 0)
    try:
        fit = e.max_resident_workers()
        assert fit >= 148
    finally:
        e.close()
    big = _engine(2, 4, 0)
    try:
        big.W = 10 ** 6                                       # pretend: the check is on the host
        with pytest.raises(native.NativeError, match="resident"):
            big.step(num_clocks=1)
    finally:
        big.W = 4
        big.close()
