"""Fused pull/push/apply protocol on virtual ranks, and the CUDA backend of the event-driven engine."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.config import PSConfig
from psb200.data import CSRBatch, ReviewGenerator
from psb200.engine.sim import NumpyBackend, PSSimulation
from psb200.models import LogRegLayout
from psb200.oracle import LogisticRegressionTask
from psb200.utils.logs import parse_log

pytestmark = pytest.mark.gpu


def _half_batch(batch):
    return CSRBatch.from_dense(batch.to_dense(np.float16).astype(np.float32), batch.labels)


def _oracle_rounds(cfg, per_worker, clocks, group):
    """Reference for ranks stepped one after the other: the `group` workers of a rank solve from the same
    weights and their summed delta is applied (w += sum/N) before the next group starts.  Sequential
    consistency is group == all workers (one snapshot per clock); the async modes on a virtual cluster
    are group == workers_per_rank (a later rank already sees the earlier ranks' pushes of that clock)."""
    lay = LogRegLayout(cfg.num_features, cfg.num_classes)
    W = lay.zeros()
    n = len(per_worker)
    losses = []
    for _ in range(clocks):
        ls = []
        for g0 in range(0, n, group):
            total = np.zeros_like(W, dtype=np.float64)
            for b in per_worker[g0:g0 + group]:
                t = LogisticRegressionTask(cfg.num_features, cfg.num_classes, None, cfg.local_solver, cfg.max_iter,
                                           cfg.sgd_lr, cfg.standardize, cfg.center)
                t.set_weights(W)
                total += t.calculate_gradients(b, eval_test=False)
                ls.append(t.loss)
            W = (W + total / n).astype(np.float32)
        losses.append(ls)
    return W, np.asarray(losses)


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("model,ranks,wpr", [(0, 1, 2), (0, 4, 2), (0, 3, 1), (-1, 2, 2), (3, 4, 1)])
def test_virtual_cluster_matches_lockstep_oracle(impl, model, ranks, wpr):
    from psb200.engine.device_engine import VirtualCluster
    F, cap, n = 256, 128, 96
    cfg = PSConfig(num_workers=ranks * wpr, num_features=F, consistency_model=model, max_buffer_size=cap,
                   min_buffer_size=16, local_solver="sgd", sgd_lr=0.8)
    gen = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0)
    full = _half_batch(gen.train_split(n * ranks * wpr))
    per_worker = [full.slice(np.arange(i * n, (i + 1) * n)) for i in range(ranks * wpr)]
    vc = VirtualCluster(cfg, ranks, wpr, impl)
    try:
        for r, e in enumerate(vc.engines):
            rows = torch.stack([torch.from_numpy(per_worker[r * wpr + j].to_dense(np.float16)) for j in range(wpr)])
            labs = torch.stack([torch.from_numpy(per_worker[r * wpr + j].labels) for j in range(wpr)])
            e.fill(rows.cuda(), labs.cuda())
        clocks = 4
        losses = []
        for _ in range(clocks):
            vc.step()
            losses.append(np.concatenate([e.read_loss().numpy().copy() for e in vc.engines]))
        torch.cuda.synchronize()
        assert all(e.status() == 0 for e in vc.engines)
        Wref, lref = _oracle_rounds(cfg, per_worker, clocks, ranks * wpr if model == 0 else wpr)
        assert np.allclose(np.asarray(losses), lref, rtol=3e-3, atol=3e-3)
        if model == 0:
            for e in vc.engines:                               # every replica holds the same snapshot
                Wd = e.weights().cpu().numpy()
                assert np.abs(Wd - Wref).max() < 3e-3 * max(1e-2, np.abs(Wref).max())
        else:
            shards = []
            for e in vc.engines:
                from psb200.ops import native
                ptr = native.lib().psb_comm_shard_weights(e.block, F, ranks)
                from psb200.engine.device_engine import _view
                L4 = (e.nv // 4 + ranks - 1) // ranks * 4
                b = min(e.rank * L4, e.nv)
                shards.append(_view(ptr, L4, e.device).cpu().numpy()[: max(0, min(L4, e.nv - b))])
            Wd = np.concatenate(shards)[: vc.engines[0].nv].reshape(F + 1, 8)
            assert np.abs(Wd - Wref).max() < 3e-3 * max(1e-2, np.abs(Wref).max())
    finally:
        vc.close()


def test_cuda_backend_tracks_numpy_backend():
    from psb200.engine.cuda_backend import CudaBackend
    gen = ReviewGenerator(mean_len=300.0)
    tr, te = _half_batch(gen.train_split(1200)), _half_batch(gen.test_split(800))
    cfg = PSConfig(num_workers=2, producer_time_per_event=0, max_clock=6, max_buffer_size=256)
    logs = []
    for mk in (lambda: NumpyBackend(cfg, te), lambda: CudaBackend(cfg, te)):
        sim = PSSimulation(cfg, tr, te, backend=mk(), server_start_ms=0.0)
        sim.run()
        logs.append(parse_log(sim.server_log.rows, "server"))
    assert logs[0].shape == logs[1].shape == (6, 6)
    assert np.abs(logs[0][:, 4] - logs[1][:, 4]).max() < 0.01          # F1 within a point at every clock
    assert np.abs(logs[0][:, 5] - logs[1][:, 5]).max() < 0.01


def test_missing_peer_raises_timeout_instead_of_hanging():
    """A rank whose peer never pushes: the in-kernel wait gives up after its cycle budget and the host sees
    PSB_E_TIMEOUT with a readable reason (failure detection)."""
    from psb200.engine.device_engine import DeviceEngine
    from psb200.ops import native
    cfg = PSConfig(num_workers=2, num_features=256, max_buffer_size=128, min_buffer_size=16)
    a = DeviceEngine(cfg, 1, 0, 2, impl=0, spin_limit_cycles=20_000_000)       # ~10 ms budget
    b = DeviceEngine(cfg, 1, 1, 2, impl=0)
    a.set_peers([a.block, b.block]), b.set_peers([a.block, b.block])
    rows = torch.zeros((1, 128, 256), dtype=torch.float16, device="cuda")
    rows[:, :, 0] = 1.0
    a.fill(rows, torch.ones((1, 128), dtype=torch.int32, device="cuda"))
    a.step()                      # clock 0: pushes, nobody else does
    a.step()                      # clock 1: waits for weights that never come
    with pytest.raises(native.NativeError, match="never arrived"):
        a.check()
    a.close(), b.close()


def test_packed_ingest_equals_dense_ingest():
    """ingest() with bitmap-packed rows leaves rings, labels and windows identical to the dense fp16 path,
    through two pipelined batches (both staging sets)."""
    from psb200.engine.device_engine import DeviceEngine
    from psb200.ops import native
    cfg = PSConfig(num_workers=3, num_features=256, max_buffer_size=128, min_buffer_size=16)
    rng = np.random.default_rng(9)
    engines = [DeviceEngine(cfg, 3, impl=0, stage_rows=96) for _ in range(2)]
    for it in range(3):
        n = 96 - 7 * it
        rows = torch.from_numpy((rng.standard_normal((n, 256)) * (rng.random((n, 256)) < 0.4)).astype(np.float16)).pin_memory()
        lab = torch.from_numpy(rng.integers(1, 6, n).astype(np.int32)).pin_memory()
        workers = rng.integers(0, 3, n).astype(np.int32)
        engines[0].ingest(rows, lab, workers)
        engines[1].ingest(native.PackedRows(96, 256, palette=bool(it & 1)).pack_dense(rows, lab), None, workers)
    torch.cuda.synchronize()
    a, b = engines
    assert torch.equal(a.X.view(torch.int16) * (a.X != 0), b.X.view(torch.int16) * (b.X != 0))
    assert torch.equal(a.labels, b.labels) and torch.equal(a.win, b.win)
    assert b.h2d_bytes < a.h2d_bytes
    a.close(), b.close()


def test_pipelined_ingest_three_sets_in_order():
    """Three batches begun ahead of their commits (one per staging set) land in the rings in begin order;
    a fourth begin without a commit is refused; CSR, dense and packed batches can be mixed."""
    from psb200.data.synthetic import random_csr
    from psb200.engine.device_engine import DeviceEngine
    from psb200.ops import native
    cfg = PSConfig(num_workers=2, num_features=256, max_buffer_size=128, min_buffer_size=16)
    rng = np.random.default_rng(2)
    piped = DeviceEngine(cfg, 2, impl=0, stage_rows=64, stage_sets=3)
    plain = DeviceEngine(cfg, 2, impl=0, stage_rows=64, stage_sets=2)
    batches = []
    for it in range(7):
        n = 64 - it
        rows = torch.from_numpy((rng.standard_normal((n, 256)) * (rng.random((n, 256)) < 0.3)).astype(np.float16)).pin_memory()
        lab = torch.from_numpy(rng.integers(1, 6, n).astype(np.int32)).pin_memory()
        batches.append((rows, lab, rng.integers(0, 2, n).astype(np.int32)))
    for rows, lab, wk in batches:
        plain.ingest(rows, lab, wk)
    csr = random_csr(20, 256, 25, seed=8)
    wk_csr = rng.integers(0, 2, 20).astype(np.int32)
    plain.ingest_sparse(csr, wk_csr)

    for k in range(3):
        rows, lab, wk = batches[k]
        piped.ingest_begin(native.PackedRows(64, 256).pack_dense(rows, lab) if k == 1 else rows, lab, wk)
    with pytest.raises(native.NativeError, match="uncommitted"):
        piped.ingest_begin(*batches[3])
    for k in range(3, 7):
        piped.ingest_commit()
        piped.ingest_begin(*batches[k])
    for _ in range(4):
        piped.ingest_commit()            # the fourth is a no-op
    piped.ingest_sparse(csr, wk_csr)
    torch.cuda.synchronize()
    assert torch.equal(piped.X.view(torch.int16) * (piped.X != 0), plain.X.view(torch.int16) * (plain.X != 0))
    assert torch.equal(piped.labels, plain.labels) and torch.equal(piped.win, plain.win)
    piped.close(), plain.close()


def test_pipelined_loss_read_returns_the_step_it_was_issued_behind():
    """read_loss_begin / read_loss_end: the copy is ordered behind the step it follows, so its ticket still holds
    that step's losses after the next step has been queued."""
    from psb200.engine.device_engine import DeviceEngine
    cfg = PSConfig(num_workers=2, num_features=256, max_buffer_size=128, min_buffer_size=16)
    rng = np.random.default_rng(4)
    eng = DeviceEngine(cfg, 2, impl=0)
    rows = torch.from_numpy((rng.standard_normal((2, 128, 256)) * (rng.random((2, 128, 256)) < 0.3)).astype(np.float16)).cuda()
    labs = torch.from_numpy(rng.integers(1, 6, (2, 128)).astype(np.int32)).cuda()
    eng.fill(rows, labs)
    eng.step()
    t1 = eng.read_loss_begin()
    first = eng.read_loss().clone()               # synchronous read of the same step
    eng.step()
    t2 = eng.read_loss_begin()
    assert t1 != t2
    assert torch.equal(eng.read_loss_end(t1), first)
    second = eng.read_loss().clone()
    assert torch.equal(eng.read_loss_end(t2), second)
    assert not torch.equal(first, second)         # the two clocks really differ
    eng.close()


@pytest.mark.parametrize("impl", [0, 1])
def test_several_sequential_clocks_in_one_launch_equal_one_launch_per_clock(impl):
    """PSB_COMM_SEQUENTIAL with num_clocks > 1: the worker CTAs stay on the GPU and meet at every clock's version
    words instead of at a kernel boundary; same weights and losses as one launch per clock."""
    from psb200.engine.device_engine import DeviceEngine
    F, cap, n, W, clocks = 256, 128, 96, 4, 5
    cfg = PSConfig(num_workers=W, num_features=F, max_buffer_size=cap, min_buffer_size=16)
    b = _half_batch(ReviewGenerator(F, 5, vocab=3000, mean_len=120.0).train_split(W * n))
    rows = torch.from_numpy(b.to_dense(np.float16)).view(W, n, F).cuda()
    labs = torch.from_numpy(b.labels.astype(np.int32)).view(W, n).cuda()
    out = []
    for per_launch in (1, clocks):
        e = DeviceEngine(cfg, W, impl=impl, spin_limit_cycles=4_000_000_000)
        e.fill(rows, labs)
        for _ in range(clocks // per_launch):
            e.step(num_clocks=per_launch)
        out.append((e.weights().cpu().numpy().copy(), e.read_loss().numpy().copy(), e.clock, e.status()))
        e.close()
    assert out[0][2] == out[1][2] == clocks and out[0][3] == out[1][3] == 0
    assert np.abs(out[0][0] - out[1][0]).max() < 2e-5 * max(1.0, np.abs(out[0][0]).max())
    assert np.allclose(out[0][1], out[1][1], rtol=2e-5, atol=2e-6)
