"""sm_100a kernels vs plain fp64/fp32 references (run on the B200 box: pytest -m gpu)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.data import CSRBatch, ReviewGenerator, random_csr
from psb200.models import LogRegLayout
from psb200.oracle import LogisticRegressionTask
from psb200.utils.metrics import evaluate as evaluate_np

pytestmark = pytest.mark.gpu

IMPLS = [0, 1]          # 0 = SIMT fp32, 1 = tcgen05/TMA


def _native():
    from psb200.ops import native
    native.lib()
    return native


def _half_batch(batch: CSRBatch) -> CSRBatch:
    """The batch as the device sees it (values rounded to fp16)."""
    d = batch.to_dense(np.float16).astype(np.float32)
    return CSRBatch.from_dense(d, batch.labels)


def _dense_data(n, F, density, seed):
    rng = np.random.default_rng(seed)
    X = (rng.random((n, F)) < density) * rng.standard_normal((n, F))
    X /= np.maximum(np.linalg.norm(X, axis=1, keepdims=True), 1e-6)
    y = rng.integers(1, 6, size=n).astype(np.int32)
    return CSRBatch.from_dense(X.astype(np.float16).astype(np.float32), y)


def _run_step(native, batches, W0, F, cap, solver, impl, standardize=True, center=True, max_iter=2, heads=None):
    dev = torch.device("cuda")
    W = len(batches)
    lay = LogRegLayout(F, 5)
    X = torch.zeros((W, cap, F), dtype=torch.float16, device=dev)
    labels = torch.zeros((W, cap), dtype=torch.int32, device=dev)
    win = torch.zeros((W, 2), dtype=torch.int32)
    for w, b in enumerate(batches):
        n = len(b)
        start = 0 if heads is None else heads[w]
        slots = (start + np.arange(n)) % cap
        X[w, torch.from_numpy(slots).to(dev)] = torch.from_numpy(b.to_dense(np.float16)).to(dev)
        labels[w, torch.from_numpy(slots).to(dev)] = torch.from_numpy(b.labels).to(dev)
        win[w, 0], win[w, 1] = (start + n) % cap, n
    win = win.to(dev)
    w_in = torch.from_numpy(W0.reshape(-1)).to(dev)
    delta = torch.zeros((W, lay.device_size), dtype=torch.float32, device=dev)
    loss = torch.zeros(W, dtype=torch.float32, device=dev)
    info = torch.zeros((W, 8), dtype=torch.int32, device=dev)
    scratch = torch.zeros((W, native.scratch_floats(F)), dtype=torch.float32, device=dev)
    sv = native.make_solver(F, lay.R, cap, solver, max_iter, standardize, center, 0.5)
    native.worker_step(sv, native.comm_none(), W, impl, X, labels, win, w_in, delta, loss, info, scratch)
    torch.cuda.synchronize()
    return delta.cpu().numpy().reshape(W, F + 1, 8), loss.cpu().numpy(), info.cpu().numpy()


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("F,cap,n", [(256, 128, 100), (1024, 256, 256)])
def test_sgd_step_matches_oracle(impl, F, cap, n):
    native = _native()
    lay = LogRegLayout(F, 5)
    batches = [_dense_data(n, F, 0.3, s) for s in range(3)]
    W0 = lay.random(1, 0.05)
    delta, loss, info = _run_step(native, batches, W0, F, cap, "sgd", impl, heads=[0, 7, cap - 5])
    for w, b in enumerate(batches):
        t = LogisticRegressionTask(F, 5, None, "sgd")
        t.set_weights(W0)
        d = t.calculate_gradients(b, eval_test=False)
        assert abs(loss[w] - t.loss) < 2e-3 * max(1.0, abs(t.loss))
        assert np.abs(delta[w] - d).max() < 2e-3 * max(1e-3, np.abs(d).max())
        assert info[w, 3] == n and not delta[w][:, 6:].any()


@pytest.mark.parametrize("impl", IMPLS)
def test_lbfgs2_step_matches_oracle(impl):
    native = _native()
    F, cap, n = 1024, 512, 384
    lay = LogRegLayout(F, 5)
    gen = ReviewGenerator(F, 5, mean_len=400.0)
    full = _half_batch(gen.train_split(2 * n))
    batches = [full.slice(np.arange(n)), full.slice(np.arange(n, 2 * n))]
    for W0 in (lay.zeros(), lay.random(3, 0.02)):
        delta, loss, info = _run_step(native, batches, W0, F, cap, "lbfgs2", impl)
        for w, b in enumerate(batches):
            t = LogisticRegressionTask(F, 5, None, "lbfgs2")
            t.set_weights(W0)
            d = t.calculate_gradients(b, eval_test=False)
            assert info[w, 1] == t.last_info.iterations == 2
            assert info[w, 0] == t.last_info.num_evals
            assert abs(loss[w] - t.loss) < 5e-3 * max(1.0, abs(t.loss))
            assert np.abs(delta[w] - d).max() < 2e-2 * np.abs(d).max()


@pytest.mark.parametrize("impl", IMPLS)
def test_no_standardize_no_center(impl):
    native = _native()
    F, cap, n = 256, 128, 128
    lay = LogRegLayout(F, 5)
    b = _dense_data(n, F, 0.4, 5)
    W0 = lay.random(2, 0.05)
    delta, loss, info = _run_step(native, [b], W0, F, cap, "lbfgs2", impl, standardize=False, center=False)
    t = LogisticRegressionTask(F, 5, None, "lbfgs2", standardize=False, center=False)
    t.set_weights(W0)
    d = t.calculate_gradients(b, eval_test=False)
    assert abs(loss[0] - t.loss) < 5e-3
    assert np.abs(delta[0] - d).max() < 2e-2 * np.abs(d).max()


@pytest.mark.parametrize("impl", IMPLS)
def test_eval_matches_reference(impl):
    native = _native()
    dev = torch.device("cuda")
    F = 1024
    lay = LogRegLayout(F, 5)
    gen = ReviewGenerator(F, 5, mean_len=300.0)
    te = _half_batch(gen.test_split(4877))
    Wt = lay.random(4, 0.5)
    X = torch.from_numpy(te.to_dense(np.float16)).to(dev)
    y = torch.from_numpy(te.labels).to(dev)
    conf = torch.zeros(64, dtype=torch.int32, device=dev)
    out = torch.zeros(2, dtype=torch.float32, device=dev)
    native.evaluate(X, y, len(te), torch.from_numpy(Wt.reshape(-1)).to(dev), F, lay.R, conf, out, impl)
    torch.cuda.synchronize()
    z = te.to_dense(np.float64) @ Wt[:F, :6].astype(np.float64) + Wt[F, :6]
    srt = np.sort(z, axis=1)
    clear = (srt[:, -1] - srt[:, -2]) > 1e-3                    # rows whose argmax is not a near-tie
    pred = np.argmax(z, axis=1)
    ref = evaluate_np(te.labels, pred, 6)
    f1, acc = out.cpu().tolist()
    n_amb = int((~clear).sum())
    assert abs(acc - ref.accuracy) <= (n_amb + 1) / len(te)
    assert abs(f1 - ref.f1) <= 2 * (n_amb + 1) / len(te) + 1e-4
    assert int(conf.sum().item()) == len(te)


def test_ring_scatter_and_store():
    native = _native()
    dev = torch.device("cuda")
    F, cap, W = 256, 64, 3
    b = random_csr(40, F, 30, seed=2)
    stride = 4 + 2 * 32
    stage = np.zeros((40, stride), dtype=np.uint32)
    rng = np.random.default_rng(0)
    ws = np.stack([rng.integers(0, W, 40), rng.permutation(cap)[:40]], axis=1).astype(np.int32)
    for i in range(40):
        idx, val, lab = b.row(i)
        stage[i, 0], stage[i, 1], stage[i, 2], stage[i, 3] = len(idx), lab, ws[i, 0], ws[i, 1]
        stage[i, 4:4 + 2 * len(idx):2] = idx.view(np.uint32)
        stage[i, 5:5 + 2 * len(idx):2] = val.view(np.uint32)
    X = torch.full((W, cap, F), 7.0, dtype=torch.float16, device=dev)
    labels = torch.zeros((W, cap), dtype=torch.int32, device=dev)
    native.ring_scatter(torch.from_numpy(stage.view(np.int32)).to(dev), 40, stride, X, labels, cap, F)
    X2 = torch.zeros_like(X)
    l2 = torch.zeros_like(labels)
    dense = torch.from_numpy(b.to_dense(np.float32)).to(dev)
    native.ring_store_dense(dense, torch.from_numpy(b.labels).to(dev), torch.from_numpy(ws).to(dev), 40, X2, l2, cap, F)
    X3 = torch.zeros_like(X)
    native.ring_store_dense(dense.half(), torch.from_numpy(b.labels).to(dev), torch.from_numpy(ws).to(dev), 40, X3, l2,
                            cap, F)
    torch.cuda.synchronize()
    ref = b.to_dense(np.float16)
    for i in range(40):
        for Xt in (X, X2, X3):
            assert np.array_equal(Xt[ws[i, 0], ws[i, 1]].cpu().numpy(), ref[i])
        assert labels[ws[i, 0], ws[i, 1]].item() == b.labels[i] == l2[ws[i, 0], ws[i, 1]].item()


@pytest.mark.parametrize("F,density", [(1024, 0.45), (1024, 0.02), (256, 1.0), (2048, 0.3), (96, 0.5)])
def test_ring_store_bitmap_matches_dense_store(F, density):
    """Bitmap-packed rows expand to exactly the fp16 ring contents the dense store writes."""
    native = _native()
    dev = torch.device("cuda")
    cap, W, n = 64, 5, 150
    rng = np.random.default_rng(F)
    rows = (rng.standard_normal((n, F)) * (rng.random((n, F)) < density)).astype(np.float16)
    rows[11] = 0
    labels_in = rng.integers(1, 6, n).astype(np.int32)
    slots = rng.permutation(W * cap)[:n]
    ws = np.stack([slots // cap, slots % cap], axis=1).astype(np.int32)
    pk = native.PackedRows(n + 3, F).pack_dense(rows, labels_in)
    X = torch.full((W, cap, F), 7.0, dtype=torch.float16, device=dev)       # stale slot contents must be overwritten
    lab = torch.zeros((W, cap), dtype=torch.int32, device=dev)
    native.ring_store_bitmap(pk.bitmap.to(dev), pk.row_off.to(dev), pk.values[:max(pk.nnz, 1)].to(dev), pk.labels.to(dev),
                             torch.from_numpy(ws).to(dev), n, X, lab, cap, F)
    X2 = torch.full_like(X, 7.0)
    lab2 = torch.zeros_like(lab)
    native.ring_store_dense(torch.from_numpy(rows).to(dev), torch.from_numpy(labels_in).to(dev),
                            torch.from_numpy(ws).to(dev), n, X2, lab2, cap, F)
    torch.cuda.synchronize()
    assert torch.equal(X.view(torch.int16), X2.view(torch.int16).where(X2 != 0, torch.zeros_like(X2).view(torch.int16)))
    assert torch.equal(lab, lab2)


@pytest.mark.parametrize("F,density,raw_rows", [(1024, 0.45, 0), (1024, 0.05, 3), (256, 1.0, 1), (2048, 0.3, 2), (96, 0.5, 5)])
def test_ring_store_palette_matches_dense_store(F, density, raw_rows):
    """Palette-packed rows (palette and raw-fallback rows mixed) expand to exactly the dense store's ring contents."""
    native = _native()
    dev = torch.device("cuda")
    cap, W, n = 64, 5, 150
    rng = np.random.default_rng(F + raw_rows)
    cnt = rng.integers(1, 6, (n, F)) * (rng.integers(0, 2, (n, F)) * 2 - 1) * (rng.random((n, F)) < density)
    rows = (cnt / np.maximum(np.linalg.norm(cnt, axis=1, keepdims=True), 1e-6)).astype(np.float16)
    rows[11] = 0
    for r in range(raw_rows):
        rows[20 + r] = (rng.standard_normal(F) * (rng.random(F) < max(density, 0.3))).astype(np.float16)
    labels_in = rng.integers(1, 6, n).astype(np.int32)
    slots = rng.permutation(W * cap)[:n]
    ws = np.stack([slots // cap, slots % cap], axis=1).astype(np.int32)
    pk = native.PackedRows(n + 3, F, palette=True).pack_dense(rows, labels_in)
    X = torch.full((W, cap, F), 7.0, dtype=torch.float16, device=dev)
    lab = torch.zeros((W, cap), dtype=torch.int32, device=dev)
    native.ring_store_palette(pk.bitmap.to(dev), pk.row_off.to(dev), pk.payload[:pk.payload_bytes + 2].to(dev), pk.labels.to(dev),
                              torch.from_numpy(ws).to(dev), n, X, lab, cap, F)
    X2 = torch.full_like(X, 7.0)
    lab2 = torch.zeros_like(lab)
    native.ring_store_dense(torch.from_numpy(rows).to(dev), torch.from_numpy(labels_in).to(dev),
                            torch.from_numpy(ws).to(dev), n, X2, lab2, cap, F)
    torch.cuda.synchronize()
    assert torch.equal(X.view(torch.int16), X2.view(torch.int16).where(X2 != 0, torch.zeros_like(X2).view(torch.int16)))
    assert torch.equal(lab, lab2)


def test_apply_delta_and_q1_compat():
    native = _native()
    dev = torch.device("cuda")
    lay = LogRegLayout(64, 5)
    w = torch.from_numpy(lay.random(1).reshape(-1)).to(dev)
    d = torch.from_numpy(lay.random(2).reshape(-1)).to(dev)
    ref = (w + 0.25 * d).cpu().numpy()
    w1 = w.clone()
    native.apply_delta(w1, d, 0.25, 64, -1)
    w2 = w.clone()
    native.apply_delta(w2, d, 0.25, 64, 5)
    torch.cuda.synchronize()
    assert np.allclose(w1.cpu().numpy(), ref, atol=1e-7)
    r2 = ref.copy()
    r2[64 * 8 + 5] = w.cpu().numpy()[64 * 8 + 5]
    assert np.allclose(w2.cpu().numpy(), r2, atol=1e-7)


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("scale", [5.0, 20.0])
def test_lbfgs2_many_line_search_trials(impl, scale):
    """Far-off start weights make the strong-Wolfe search take several trials per iteration (6 and 10 loss
    evaluations per solve in the oracle); the tcgen05 kernel serves all but the first trial of an iteration
    from cached logits and must still walk the oracle's path."""
    native = _native()
    F, cap, n = 1024, 128, 128
    lay = LogRegLayout(F, 5)
    gen = ReviewGenerator(F, 5, mean_len=400.0)
    b = _half_batch(gen.train_split(256)).slice(np.arange(n))
    W0 = lay.random(3, scale)
    delta, loss, info = _run_step(native, [b], W0, F, cap, "lbfgs2", impl, standardize=False, center=False)
    t = LogisticRegressionTask(F, 5, None, "lbfgs2", standardize=False, center=False)
    t.set_weights(W0)
    d = t.calculate_gradients(b, eval_test=False)
    assert t.last_info.num_evals > 4                           # the scenario this test is about
    assert info[0, 0] == t.last_info.num_evals and info[0, 1] == 2 and info[0, 2] == 0
    assert abs(loss[0] - t.loss) < 5e-3 * max(1.0, abs(t.loss))
    assert np.abs(delta[0] - d).max() < 2e-2 * np.abs(d).max()
