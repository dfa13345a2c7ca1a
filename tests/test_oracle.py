"""L-BFGS restatement and the LR objective."""
import numpy as np
import scipy.optimize

from psb200.data import ReviewGenerator
from psb200.oracle import LogisticRegressionTask, csr_matrix, feature_std, lbfgs, loss_and_grad


def test_lbfgs_converges_like_scipy_on_rosenbrock():
    f = lambda x: (scipy.optimize.rosen(x), scipy.optimize.rosen_der(x))
    res = lbfgs.minimize(f, np.array([-1.2, 1.0, -0.5, 0.8]), max_iter=200, tol=1e-12)
    assert np.allclose(res.x, 1.0, atol=1e-4)


def test_two_iterations_decrease_loss_and_count_evals():
    rng = np.random.default_rng(0)
    A = rng.standard_normal((20, 8))
    b = rng.standard_normal(20)
    f = lambda x: (0.5 * np.sum((A @ x - b) ** 2), A.T @ (A @ x - b))
    res = lbfgs.minimize(f, np.zeros(8), max_iter=2)
    assert res.iterations == 2 and len(res.loss_history) == 3
    assert res.loss_history[2] < res.loss_history[1] < res.loss_history[0]
    assert res.num_evals >= 3


def test_gradient_matches_finite_differences():
    g = ReviewGenerator(num_features=32, vocab=300, mean_len=10)
    b = g.train_split(50)
    X = csr_matrix(b)
    rng = np.random.default_rng(1)
    W = rng.standard_normal((32, 6)) * 0.1
    c = rng.standard_normal(6) * 0.1
    loss, dW, dc = loss_and_grad(W, c, X, b.labels.astype(np.int64))
    eps = 1e-6
    for (i, j) in [(0, 0), (5, 3), (31, 5)]:
        Wp = W.copy()
        Wp[i, j] += eps
        lp = loss_and_grad(Wp, c, X, b.labels.astype(np.int64))[0]
        assert abs((lp - loss) / eps - dW[i, j]) < 1e-4
    cp = c.copy()
    cp[2] += eps
    assert abs((loss_and_grad(W, cp, X, b.labels.astype(np.int64))[0] - loss) / eps - dc[2]) < 1e-4


def test_feature_std_is_unbiased_sample_std():
    g = ReviewGenerator(num_features=16, vocab=100, mean_len=8)
    b = g.train_split(30)
    assert np.allclose(feature_std(csr_matrix(b)), b.to_dense(np.float64).std(axis=0, ddof=1))


def test_task_delta_semantics_and_zero_variance_features():
    g = ReviewGenerator(num_features=64, vocab=400, mean_len=10)
    b = g.train_split(64)
    te = g.test_split(200)
    t = LogisticRegressionTask(64, 5, te)
    t.initialize(True)
    assert not t.get_weights().any()                      # zeros despite "randomly" (Q3)
    W0 = t.layout.random(1, 0.05)
    t.set_weights(W0)
    d = t.calculate_gradients(b)
    assert d.shape == (65, 8) and not d[:, 6:].any()
    dead = np.where(b.to_dense().std(axis=0, ddof=1) == 0)[0]
    if len(dead):                                          # Q5: delta = -w for features absent from the buffer
        assert np.allclose((W0 + d)[dead, :6], 0.0, atol=1e-7)
    assert abs((W0 + d)[:64, :6].mean(axis=1)).max() < 1e-6   # centred across classes
    assert t.loss < np.log(6) and 0 <= t.metrics.f1 <= 1
