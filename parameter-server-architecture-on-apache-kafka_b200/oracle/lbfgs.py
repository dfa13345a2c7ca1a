"""L-BFGS with a strong-Wolfe line search — the local solver of the reference.

The reference calls Spark's ``LogisticRegression().setMaxIter(2).fit``
(``LogisticRegressionTaskSpark.java:179-184``) which runs Breeze's L-BFGS
(memory 10) with its strong-Wolfe cubic line search (c1=1e-4, c2=0.9, at most
10 bracketing and 10 zoom steps; first step length 1/||direction||).  None of
that code is in the reference tree; this file states the algorithm so that the
CUDA solver (``csrc/ps_worker.cu``) and the CPU oracle walk the same path.

``fun(x) -> (value, gradient)``; every call is one pass over the worker's buffer.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, List, Tuple

import numpy as np

C1 = 1e-4
C2 = 0.9
MAX_LS_ITER = 10
MAX_ZOOM_ITER = 10
MEMORY = 10


class LineSearchFailed(Exception):
    pass


@dataclass
class LBFGSResult:
    x: np.ndarray
    value: float
    grad: np.ndarray
    loss_history: List[float] = field(default_factory=list)
    num_evals: int = 0
    iterations: int = 0
    status: str = "max_iter"


def _interp(l, r):
    """Cubic interpolation between brackets (t, dd, fval) — Nocedal & Wright p.57."""
    lt, ldd, lf = l
    rt, rdd, rf = r
    d1 = ldd + rdd - 3.0 * (lf - rf) / (lt - rt)
    disc = d1 * d1 - ldd * rdd
    d2 = math.sqrt(disc) if disc > 0 else 0.0
    denom = rdd - ldd + 2.0 * d2
    t = rt - (rt - lt) * (rdd + d2 - d1) / denom if denom != 0 else 0.5 * (lt + rt)
    lb = lt + 0.1 * (rt - lt)
    ub = lt + 0.9 * (rt - lt)
    if not (t >= lb):           # also catches NaN
        return lb
    if t > ub:
        return ub
    return t


def strong_wolfe(phi: Callable[[float], Tuple[float, float]], f0: float, dd0: float, init: float) -> float:
    """phi(t) -> (value, directional derivative).  Returns the accepted step."""
    if dd0 > 0:
        raise LineSearchFailed("non-descent direction")

    def zoom(lo, hi):
        for _ in range(MAX_ZOOM_ITER):
            t = _interp(hi, lo) if lo[0] > hi[0] else _interp(lo, hi)
            fv, dd = phi(t)
            c = (t, dd, fv)
            if fv > f0 + C1 * t * dd0 or fv >= lo[2]:
                hi = c
            else:
                if abs(dd) <= C2 * abs(dd0):
                    return t
                if dd * (hi[0] - lo[0]) >= 0:
                    hi = lo
                lo = c
        raise LineSearchFailed("zoom failed")

    t = init
    low = (0.0, dd0, f0)
    for i in range(MAX_LS_ITER):
        fv, dd = phi(t)
        if math.isinf(fv) or math.isnan(fv):
            t /= 2.0
            continue
        c = (t, dd, fv)
        if fv > f0 + C1 * t * dd0 or (fv >= low[2] and i > 0):
            return zoom(low, c)
        if abs(dd) <= C2 * abs(dd0):
            return t
        if dd >= 0:
            return zoom(c, low)
        low = c
        t *= 1.5
    raise LineSearchFailed("line search failed")


def two_loop(grad: np.ndarray, s_hist: List[np.ndarray], y_hist: List[np.ndarray]) -> np.ndarray:
    """Descent direction -H*grad; histories newest first."""
    d = grad.astype(np.float64).copy()
    m = len(s_hist)
    if m:
        sy = float(s_hist[0] @ y_hist[0])
        yy = float(y_hist[0] @ y_hist[0])
        if sy < 0 or math.isnan(sy):
            raise LineSearchFailed("NaN history")
        diag = sy / yy
    else:
        diag = 1.0
    alpha = np.zeros(m)
    rho = np.zeros(m)
    for i in range(m):
        rho[i] = float(s_hist[i] @ y_hist[i])
        alpha[i] = float(s_hist[i] @ d) / rho[i]
        d -= alpha[i] * y_hist[i]
    d *= diag
    for i in range(m - 1, -1, -1):
        beta = float(y_hist[i] @ d) / rho[i]
        d += (alpha[i] - beta) * s_hist[i]
    return -d


def minimize(fun: Callable[[np.ndarray], Tuple[float, np.ndarray]], x0: np.ndarray, max_iter: int = 2,
             tol: float = 1e-6) -> LBFGSResult:
    x = np.asarray(x0, dtype=np.float64).copy()
    evals = 0

    def call(z):
        nonlocal evals
        evals += 1
        v, g = fun(z)
        return float(v), np.asarray(g, dtype=np.float64)

    value, grad = call(x)
    res = LBFGSResult(x, value, grad, [value], evals)
    s_hist: List[np.ndarray] = []
    y_hist: List[np.ndarray] = []
    initial_value = value
    for it in range(max_iter):
        gnorm = float(np.linalg.norm(grad))
        if gnorm <= max(tol * abs(value), 1e-8):
            res.status = "gradient_converged"
            break
        try:
            d = two_loop(grad, s_hist, y_hist)
            dd0 = float(grad @ d)
            cache = {}

            def phi(t):
                v, g = call(x + t * d)
                cache["last"] = (t, v, g)
                return v, float(g @ d)

            init = 1.0 / float(np.linalg.norm(d)) if it == 0 else 1.0
            alpha = strong_wolfe(phi, value, dd0, init)
            if alpha * gnorm < 1e-10:
                raise LineSearchFailed("step size underflow")
        except LineSearchFailed as e:
            res.status = f"line_search_failed: {e}"
            break
        t, v, g = cache["last"]
        assert t == alpha
        x_new = x + alpha * d
        s_hist.insert(0, x_new - x)
        y_hist.insert(0, g - grad)
        del s_hist[MEMORY:], y_hist[MEMORY:]
        x, value, grad = x_new, v, g
        res.loss_history.append(value)
        res.iterations = it + 1
        # function-value convergence (Breeze: |f - max(recent f)| <= tol * initial f needs >= 2 values)
        if it + 1 < max_iter and len(res.loss_history) > 2 and \
                abs(value - max(res.loss_history[-3:-1])) <= max(tol * abs(initial_value), 1e-8):
            res.status = "function_converged"
            break
    res.x, res.value, res.grad, res.num_evals = x, value, grad, evals
    return res
