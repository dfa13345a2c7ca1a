"""Sharded sparse path vs a dense NumPy SGD reference (virtual shards on one GPU)."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

from psb200.data import random_csr
from psb200.utils.metrics import evaluate as evaluate_np

pytestmark = pytest.mark.gpu


def _softmax_residual(X, y, W, b, R=6):
    z = X @ W[:, :R] + b[:R]
    z -= z.max(axis=1, keepdims=True)
    p = np.exp(z)
    p /= p.sum(axis=1, keepdims=True)
    loss = -np.log(p[np.arange(len(y)), y])
    p[np.arange(len(y)), y] -= 1
    return loss, p


@pytest.mark.parametrize("model,ranks", [(0, 1), (0, 4), (-1, 3)])
def test_sparse_steps_match_dense_sgd(model, ranks):
    from psb200.engine.sparse_engine import DeviceCSR, VirtualSparseCluster
    F, B, eta = 5000, 256, 0.7
    vc = VirtualSparseCluster(F, ranks, model, eta)
    try:
        W = np.zeros((F, 8))
        b = np.zeros(8)
        for step in range(3):
            batches = [random_csr(B, F, 40, seed=10 * step + r) for r in range(ranks)]
            dev = [DeviceCSR(bt, vc.engines[0].device) for bt in batches]
            vc.step(dev)
            losses = [e.mean_loss(d) for e, d in zip(vc.engines, dev)]
            # reference: sequential = all workers read the same W; eventual (virtual) = rank after rank
            dW, db = np.zeros_like(W), np.zeros_like(b)
            ref_losses = []
            for bt in batches:
                X = bt.to_dense(np.float64)
                Wr, br = (W, b) if model == 0 else (W + dW, b + db)
                loss, r = _softmax_residual(X, bt.labels.astype(np.int64), Wr, br)
                ref_losses.append(loss.mean())
                dW[:, :6] += -eta / (B * ranks) * (X.T @ r)
                db[:6] += -eta / (B * ranks) * r.sum(axis=0)
            W, b = W + dW, b + db
            assert np.allclose(losses, ref_losses, rtol=2e-4, atol=2e-4)
        Wd, bd = vc.weights()
        assert np.abs(Wd - W).max() < 2e-4 * max(1.0, np.abs(W).max()) and np.abs(bd - b).max() < 2e-4
        assert all(e.status() == 0 for e in vc.engines)
        # evaluation over the sharded model
        te = random_csr(500, F, 40, seed=99)
        f1, acc = vc.engines[0].evaluate(DeviceCSR(te, vc.engines[0].device))
        z = te.to_dense(np.float64) @ W[:, :6] + b[:6]
        ref = evaluate_np(te.labels, z.argmax(axis=1), 6)
        assert abs(acc - ref.accuracy) < 0.01 and abs(f1 - ref.f1) < 0.01
    finally:
        vc.close()


@pytest.mark.parametrize("ranks", [1, 4])
def test_slab_exchange_matches_dense_sgd(ranks):
    from psb200.engine.sparse_engine import DeviceCSR, VirtualSlabCluster
    F, B, eta = 5000, 256, 0.7
    vc = VirtualSlabCluster(F, ranks, eta, max_nnz=B * 40)
    try:
        W, b = np.zeros((F, 8)), np.zeros(8)
        for step in range(3):
            batches = [random_csr(B, F, 40, seed=50 * step + r) for r in range(ranks)]
            dev = [DeviceCSR(bt, vc.engines[0].device) for bt in batches]
            vc.step(dev)
            losses = [e.mean_loss(d) for e, d in zip(vc.engines, dev)]
            dW, db, ref_losses = np.zeros_like(W), np.zeros_like(b), []
            for bt in batches:
                X = bt.to_dense(np.float64)
                loss, r = _softmax_residual(X, bt.labels.astype(np.int64), W, b)
                ref_losses.append(loss.mean())
                dW[:, :6] += -eta / (B * ranks) * (X.T @ r)
                db[:6] += -eta / (B * ranks) * r.sum(axis=0)
            W, b = W + dW, b + db
            assert np.allclose(losses, ref_losses, rtol=2e-4, atol=2e-4)
        Wd, bd = vc.weights()
        assert np.abs(Wd - W).max() < 2e-4 * max(1.0, np.abs(W).max()) and np.abs(bd - b).max() < 2e-4
        assert all(e.status() == 0 for e in vc.engines)
    finally:
        vc.close()
