"""Branches of the CPU oracle and host logic that the end-to-end CPU tests do not reach: every exit of the strong
Wolfe line search (the device solver in csrc/solver.cuh follows the same decision tree), the L-BFGS stop reasons, the
plain-gradient local solver, the Q1 compatibility switch of the CPU engine, row-wise CSV readers, rank checkpoints."""
import math
import os

import numpy as np
import pytest

from psb200.config import PSConfig
from psb200.consistency.tracker import ClockTracker, ConsistencyScheduler, ProtocolViolation
from psb200.data import CSRBatch, ReviewGenerator, iter_rows, parse_line, read_csv, write_csv
from psb200.engine.sim import PSSimulation
from psb200.models import LogRegLayout
from psb200.oracle import LogisticRegressionTask, lbfgs
from psb200.utils import checkpoint


def _phi(f, df):
    calls = []

    def phi(t):
        calls.append(t)
        return f(t), df(t)
    return phi, calls


def _is_strong_wolfe(f, df, t):
    f0, d0 = f(0.0), df(0.0)
    return f(t) <= f0 + lbfgs.C1 * t * d0 + 1e-12 and abs(df(t)) <= lbfgs.C2 * abs(d0) + 1e-12


@pytest.mark.parametrize("name,f,df,init", [
    # accepted at the first trial step
    ("first", lambda t: (t - 1.0) ** 2, lambda t: 2 * (t - 1.0), 1.0),
    # first step overshoots the Armijo line: zoom between 0 and the trial, upper end moves in
    ("overshoot", lambda t: (t - 1.0) ** 2, lambda t: 2 * (t - 1.0), 40.0),
    # first step lands where the slope is already positive but the value is fine: zoom(c, low)
    ("positive_slope", lambda t: (t - 1.0) ** 4, lambda t: 4 * (t - 1.0) ** 3, 1.9),
    # the step has to be extended (x1.5) several times before the curvature condition holds
    ("extend", lambda t: (t - 30.0) ** 2, lambda t: 2 * (t - 30.0), 0.5),
    # a wiggly valley: the zoom moves its lower end and swaps the bracket
    ("wiggle", lambda t: (t - 2.0) ** 2 + 0.3 * math.sin(9 * t), lambda t: 2 * (t - 2.0) + 2.7 * math.cos(9 * t), 6.0),
])
def test_strong_wolfe_returns_a_strong_wolfe_step_on_every_path(name, f, df, init):
    if df(0.0) >= 0:
        pytest.skip("not a descent direction at 0")
    phi, calls = _phi(f, df)
    t = lbfgs.strong_wolfe(phi, f(0.0), df(0.0), init)
    assert t > 0 and _is_strong_wolfe(f, df, t), (name, t)
    if name == "first":
        assert calls == [1.0]
    if name == "extend":
        assert calls[:3] == [0.5, 0.75, 1.125]


def test_strong_wolfe_halves_the_step_on_inf_or_nan_and_rejects_non_descent():
    f = lambda t: (t - 1.0) ** 2 if t < 3.0 else (math.inf if t < 6.0 else math.nan)
    df = lambda t: 2 * (t - 1.0)
    phi, calls = _phi(f, df)
    t = lbfgs.strong_wolfe(phi, f(0.0), df(0.0), 8.0)
    assert calls[:3] == [8.0, 4.0, 2.0] and _is_strong_wolfe(f, df, t)
    with pytest.raises(lbfgs.LineSearchFailed, match="non-descent"):
        lbfgs.strong_wolfe(phi, 0.0, +1.0, 1.0)


def test_strong_wolfe_gives_up_with_a_reason():
    # the value always exceeds the Armijo line although the slope at 0 says descent: the zoom can never succeed
    phi, _ = _phi(lambda t: 1.0 + t, lambda t: -1.0)
    with pytest.raises(lbfgs.LineSearchFailed, match="zoom failed"):
        lbfgs.strong_wolfe(phi, 1.0, -1.0, 1.0)
    # never finite: the step is halved until the iteration budget is gone
    phi, calls = _phi(lambda t: math.nan, lambda t: -1.0)
    with pytest.raises(lbfgs.LineSearchFailed, match="line search failed"):
        lbfgs.strong_wolfe(phi, 1.0, -1.0, 1.0)
    assert len(calls) == lbfgs.MAX_LS_ITER
    # unbounded below along the ray: the step is extended until the budget is gone
    phi, calls = _phi(lambda t: -t, lambda t: -1.0)
    with pytest.raises(lbfgs.LineSearchFailed, match="line search failed"):
        lbfgs.strong_wolfe(phi, 0.0, -1.0, 1.0)


def test_cubic_interpolation_is_clamped_into_the_bracket():
    assert lbfgs._interp((0.0, -1.0, 1.0), (1.0, -0.1, 0.3)) == 0.9         # still falling at the far end: 90 % clamp
    assert lbfgs._interp((0.0, -1.0, 1.0), (1.0, -1.0, 0.0)) == 0.5         # a straight line has no cubic minimum: midpoint
    t = lbfgs._interp((0.0, -2.0, 1.0), (2.0, 2.0, 1.0))             # symmetric parabola: its minimum
    assert abs(t - 1.0) < 1e-12
    assert lbfgs._interp((0.0, math.nan, 1.0), (1.0, 1.0, 0.0)) == 0.1       # NaN falls to the lower clamp


def test_lbfgs_stop_reasons_and_agreement_with_scipy():
    from scipy.optimize import minimize as sp_minimize
    A = np.diag([1.0, 10.0, 100.0])
    b = np.array([1.0, -2.0, 3.0])
    fun = lambda x: (0.5 * x @ A @ x - b @ x, A @ x - b)
    x_star = np.linalg.solve(A, b)
    res = lbfgs.minimize(fun, np.zeros(3), max_iter=100, tol=1e-12)
    assert np.allclose(res.x, x_star, atol=1e-6) and res.iterations < 100
    assert res.status in ("gradient_converged", "function_converged") and res.loss_history[-1] <= res.loss_history[0]
    ref = sp_minimize(lambda x: fun(x)[0], np.zeros(3), jac=lambda x: fun(x)[1], method="L-BFGS-B")
    assert abs(res.value - ref.fun) < 1e-8
    res = lbfgs.minimize(fun, x_star, max_iter=5)                     # starts at the optimum
    assert res.status == "gradient_converged" and res.iterations == 0 and res.num_evals == 1
    res = lbfgs.minimize(fun, np.zeros(3), max_iter=2)                # the reference's two iterations
    assert res.status == "max_iter" and res.iterations == 2 and len(res.loss_history) == 3
    # a direction along which the function never gets finite: reported, not raised; x stays where it was
    bad = lambda x: (0.0, np.ones(3)) if not x.any() else (math.nan, np.ones(3))
    res = lbfgs.minimize(bad, np.zeros(3), max_iter=3)
    assert res.status.startswith("line_search_failed") and not res.x.any() and res.iterations == 0
    with pytest.raises(lbfgs.LineSearchFailed, match="NaN history"):
        lbfgs.two_loop(np.ones(2), [np.array([1.0, 0.0])], [np.array([-1.0, 0.0])])      # s.y < 0


def _buffer(n=80, F=32, seed=0):
    return ReviewGenerator(num_features=F, vocab=300, mean_len=10, seed=seed).train_split(n)


def test_sgd_local_solver_is_one_gradient_step_and_unknown_solvers_are_rejected():
    buf = _buffer()
    t = LogisticRegressionTask(32, 5, local_solver="sgd", sgd_lr=0.25, standardize=False)
    delta = t.calculate_gradients(buf, eval_test=False)
    X = buf.to_dense(np.float64)
    R = t.R
    p = np.full((len(buf), R), 1.0 / R)                                # softmax of zero logits
    p[np.arange(len(buf)), buf.labels] -= 1.0
    gW, gb = X.T @ p / len(buf), p.mean(axis=0)
    assert np.allclose(delta[:32, :R], -0.25 * gW, atol=1e-6) and np.allclose(delta[32, :R], -0.25 * gb, atol=1e-6)
    assert t.last_info.status == "sgd" and abs(t.loss - math.log(R)) < 1e-9
    with pytest.raises(ValueError, match="unknown local solver"):
        LogisticRegressionTask(32, 5, local_solver="adam").calculate_gradients(buf, eval_test=False)
    empty = CSRBatch(np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32), np.zeros(0, np.int32), 32)
    with pytest.raises(ValueError, match="no data"):
        t.calculate_gradients(empty, eval_test=False)


def test_standardisation_switch_changes_the_fit_but_not_the_starting_loss():
    buf = _buffer(120)
    a = LogisticRegressionTask(32, 5, standardize=True)
    b = LogisticRegressionTask(32, 5, standardize=False)
    da, db = a.calculate_gradients(buf, eval_test=False), b.calculate_gradients(buf, eval_test=False)
    assert a.last_info.loss_history[0] == pytest.approx(b.last_info.loss_history[0])      # log(6) from zero weights
    assert not np.allclose(da, db) and a.loss < math.log(6) and b.loss < math.log(6)


def test_cpu_engine_compat_q1_never_updates_the_last_intercept():
    g = ReviewGenerator(num_features=64, vocab=400, mean_len=12)
    tr, te = g.train_split(400), g.test_split(100)
    finals = {}
    for q1 in (False, True):
        cfg = PSConfig(num_workers=2, num_features=64, producer_time_per_event=0, max_clock=3, compat_q1=q1,
                       min_buffer_size=16, max_buffer_size=64)
        sim = PSSimulation(cfg, tr, te, server_start_ms=0.0)
        sim.run()
        finals[q1] = sim.backend.weights.copy()
    L = LogRegLayout(64, 5)
    assert finals[True][L.F, L.R - 1] == 0.0 and finals[False][L.F, L.R - 1] != 0.0       # key 6149's analogue (Q1)
    assert np.abs(finals[True][:L.F]).sum() > 0


def test_row_wise_csv_readers_and_the_empty_file(tmp_path):
    b = _buffer(25)
    path = os.path.join(tmp_path, "d.csv")
    write_csv(path, b)
    rows = list(iter_rows(path))
    assert len(rows) == 25
    for r, (idx, val, lab) in enumerate(rows):
        i0, v0, l0 = b.row(r)
        assert idx.tolist() == i0.tolist() and np.allclose(val, v0, atol=1e-6) and lab == l0
    idx, val, lab = parse_line("0,0.5,0,-2,3\r\n")
    assert idx.tolist() == [1, 3] and val.tolist() == [0.5, -2.0] and lab == 3
    with open(path, "w") as f:
        f.write(",".join(str(i) for i in range(32)) + ",Score\n")
    e = read_csv(path)
    assert len(e) == 0 and e.num_features == 32
    assert read_csv(path, num_features=7).num_features == 7


def test_rank_state_round_trip_and_shape_check(tmp_path):
    torch = pytest.importorskip("torch")
    state = {"clock": 12, "X": torch.arange(12, dtype=torch.float16).reshape(3, 4), "buffers": [np.arange(5.0)]}
    assert checkpoint.load_rank_state(str(tmp_path), 3) is None
    path = checkpoint.save_rank_state(str(tmp_path), 3, state)
    assert path.endswith("rank-3.pt") and not os.path.exists(path + ".tmp")
    back = checkpoint.load_rank_state(str(tmp_path), 3)
    assert back["clock"] == 12 and torch.equal(back["X"], state["X"]) and np.array_equal(back["buffers"][0], np.arange(5.0))
    L = LogRegLayout(16, 5)
    checkpoint.save(str(tmp_path), 4, L.zeros(), L)
    with pytest.raises(ValueError, match="different model shape"):
        checkpoint.load(str(tmp_path), LogRegLayout(32, 5))


def test_tracker_small_print():
    with pytest.raises(ValueError):
        ClockTracker(0)
    with pytest.raises(ValueError):
        ConsistencyScheduler(2, -2)
    assert [ConsistencyScheduler(2, m).name for m in (-1, 0, 7)] == ["eventual", "sequential", "bounded_delay_7"]
    t = ClockTracker(3)
    for w in range(3):
        t.received(w, 0)
    t.sent_all(1)
    assert all(s.sent for s in t.status)
    with pytest.raises(ProtocolViolation):
        t.received(0, 5)
