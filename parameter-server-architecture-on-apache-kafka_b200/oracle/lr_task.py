"""CPU oracle of the reference's ML task (model + local solver + evaluation).

Restates ``LogisticRegressionTaskSpark.java:56-251`` without Spark: multinomial
logistic regression over R = numClasses+1 class rows, zero initial weights
(``:98-104``), ``calculateGradients`` = run the local solver from the pulled
weights on the worker's buffer and return *new - old* per key (``:179-218``,
quirk Q2), ``calculateTestMetrics`` = weighted F1 + accuracy of the current
weights on the cached test set (``:223-251``, ``Metrics.java:15-24``).

Spark defaults reproduced behind switches: ``standardize`` (features divided by
their unbiased sample std over the buffer; zero-variance features get zero
coefficients — Q5) and ``center`` (multinomial coefficients and intercepts are
centred across classes when regParam == 0).

Weights use the device layout ``W[(F+1), 8]`` so results compare 1:1 with the
kernels; this class is also the fp64 reference of the kernel numerics tests.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np
import scipy.sparse as sp

from ..data.csr import CSRBatch
from ..models.logreg import LogRegLayout
from ..utils.metrics import Metrics, evaluate
from . import lbfgs


def csr_matrix(batch: CSRBatch) -> sp.csr_matrix:
    return sp.csr_matrix((batch.values.astype(np.float64), batch.indices, batch.indptr),
                         shape=(len(batch), batch.num_features))


def feature_std(X: sp.csr_matrix) -> np.ndarray:
    """Unbiased per-feature sample std over the rows of X (zeros included)."""
    n = X.shape[0]
    s1 = np.asarray(X.sum(axis=0)).ravel()
    s2 = np.asarray(X.multiply(X).sum(axis=0)).ravel()
    if n < 2:
        return np.zeros_like(s1)
    var = (s2 - s1 * s1 / n) / (n - 1)
    return np.sqrt(np.maximum(var, 0.0))


def loss_and_grad(Wc: np.ndarray, b: np.ndarray, X: sp.csr_matrix, y: np.ndarray) -> Tuple[float, np.ndarray, np.ndarray]:
    """Mean softmax cross-entropy; Wc[F,R], b[R]; returns (loss, dWc, db)."""
    n = X.shape[0]
    z = X @ Wc + b
    z -= z.max(axis=1, keepdims=True)
    e = np.exp(z)
    p = e / e.sum(axis=1, keepdims=True)
    loss = float(-np.log(p[np.arange(n), y]).mean())
    r = p
    r[np.arange(n), y] -= 1.0
    r /= n
    return loss, np.asarray(X.T @ r), r.sum(axis=0)


@dataclass
class SolveInfo:
    loss: float
    num_evals: int
    iterations: int
    status: str
    loss_history: list


class LogisticRegressionTask:
    def __init__(self, num_features: int = 1024, num_classes: int = 5, test: Optional[CSRBatch] = None,
                 local_solver: str = "lbfgs2", max_iter: int = 2, sgd_lr: float = 0.5,
                 standardize: bool = True, center: bool = True):
        self.layout = LogRegLayout(num_features, num_classes)
        self.F, self.R = self.layout.F, self.layout.R
        self.local_solver, self.max_iter, self.sgd_lr = local_solver, max_iter, sgd_lr
        self.standardize, self.center = standardize, center
        self.weights = self.layout.zeros()
        self.loss = 1.0                                  # LogisticRegressionTaskSpark.java:45
        self.metrics: Optional[Metrics] = None
        self.is_initialized = False
        self._test = test
        self._Xtest = csr_matrix(test) if test is not None else None
        self.last_info: Optional[SolveInfo] = None

    # -- reference API ------------------------------------------------------
    def initialize(self, randomly_initialize_weights: bool) -> None:
        if randomly_initialize_weights:
            self.weights = self.layout.zeros()           # zeros despite the name (Q3)
        self.is_initialized = True

    def set_weights(self, W: np.ndarray) -> None:
        self.weights = np.array(W, dtype=np.float32).reshape(self.layout.device_rows, self.layout.CP)

    def get_weights(self) -> np.ndarray:
        return self.weights

    def calculate_gradients(self, buffer: CSRBatch, eval_test: bool = True) -> np.ndarray:
        """Local solve from the current weights; returns delta in device layout (float32)."""
        W0 = self.weights.astype(np.float64)
        W1, info = self.solve(W0, buffer)
        self.loss = info.loss
        self.last_info = info
        W1f = W1.astype(np.float32)
        if eval_test and self._Xtest is not None:
            self.metrics = self._evaluate(W1f)
        return (W1f - self.weights).astype(np.float32)    # float32 difference of float32 casts (:195-218)

    def calculate_test_metrics(self) -> Metrics:
        self.metrics = self._evaluate(self.weights)
        return self.metrics

    # -- internals ------------------------------------------------------------
    def _evaluate(self, W: np.ndarray) -> Metrics:
        z = self._Xtest @ W[: self.F, : self.R].astype(np.float64) + W[self.F, : self.R]
        pred = np.argmax(z, axis=1)
        return evaluate(self._test.labels, pred, self.R)

    def solve(self, W0: np.ndarray, buffer: CSRBatch) -> Tuple[np.ndarray, SolveInfo]:
        F, R = self.F, self.R
        X = csr_matrix(buffer)
        y = buffer.labels.astype(np.int64)
        if len(y) == 0:
            raise ValueError("There is no data in the buffer")       # WorkerTrainingProcessor.java:131-133
        if self.standardize:
            std = feature_std(X)
            inv = np.where(std > 0, 1.0 / np.where(std > 0, std, 1.0), 0.0)
        else:
            std = np.ones(F)
            inv = np.ones(F)
        Xs = X @ sp.diags(inv) if self.standardize else X
        x0 = np.concatenate([(W0[:F, :R] * std[:, None]).ravel(), W0[F, :R]])

        def fun(x):
            Wc = x[: F * R].reshape(F, R)
            b = x[F * R:]
            loss, dW, db = loss_and_grad(Wc, b, Xs, y)
            return loss, np.concatenate([dW.ravel(), db])

        if self.local_solver == "lbfgs2":
            res = lbfgs.minimize(fun, x0, max_iter=self.max_iter)
            x1 = res.x
            info = SolveInfo(res.value, res.num_evals, res.iterations, res.status, res.loss_history)
        elif self.local_solver == "sgd":
            loss, g = fun(x0)
            x1 = x0 - self.sgd_lr * g
            info = SolveInfo(loss, 1, 1, "sgd", [loss])
        else:
            raise ValueError(f"unknown local solver {self.local_solver!r}")
        W1 = np.zeros_like(W0)
        W1[:F, :R] = x1[: F * R].reshape(F, R) * inv[:, None]
        W1[F, :R] = x1[F * R:]
        if self.center and self.local_solver == "lbfgs2":
            W1[:F, :R] -= W1[:F, :R].mean(axis=1, keepdims=True)
            W1[F, :R] -= W1[F, :R].mean()
        return W1, info
