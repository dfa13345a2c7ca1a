"""Parameter layout of the multinomial logistic-regression task.

Reference key space (``LogisticRegressionTaskSpark.java:98-140, 173, 195-218``):
keys ``0 .. R*F-1`` hold the coefficient matrix ``R x F`` (R = numClasses+1 = 6
class rows, F = 1024) in column-major order, i.e. ``key = feature*R + class``;
keys ``R*F .. R*F+R-1`` hold the intercepts.  6150 floats in total.

Device layout: ``W[(F+1), 8]`` fp32 — one 32-byte sector per feature, row ``F``
is the intercept row, class columns 6 and 7 are padding that never enters the
softmax.  One sector per non-zero is exactly what the CSR gather loads.
"""
from __future__ import annotations

import numpy as np

from ..config import CLASS_PAD


class LogRegLayout:
    def __init__(self, num_features: int = 1024, num_classes: int = 5):
        self.F = int(num_features)
        self.R = int(num_classes) + 1
        if self.R > CLASS_PAD:
            raise ValueError(f"at most {CLASS_PAD - 1} classes are supported by the 32-byte row layout")
        self.CP = CLASS_PAD
        self.num_params = self.R * self.F + self.R
        self.device_rows = self.F + 1
        self.device_size = self.device_rows * self.CP

    # -- reference keys <-> device ---------------------------------------
    def keys_to_device(self, flat: np.ndarray) -> np.ndarray:
        flat = np.asarray(flat, dtype=np.float32)
        if flat.shape != (self.num_params,):
            raise ValueError(f"expected {self.num_params} values, got {flat.shape}")
        W = np.zeros((self.device_rows, self.CP), dtype=np.float32)
        W[: self.F, : self.R] = flat[: self.R * self.F].reshape(self.F, self.R)
        W[self.F, : self.R] = flat[self.R * self.F:]
        return W

    def device_to_keys(self, W: np.ndarray) -> np.ndarray:
        W = np.asarray(W, dtype=np.float32).reshape(self.device_rows, self.CP)
        return np.concatenate([W[: self.F, : self.R].reshape(-1), W[self.F, : self.R]])

    def key_of(self, feature: int, cls: int) -> int:
        return feature * self.R + cls

    def intercept_key(self, cls: int) -> int:
        return self.R * self.F + cls

    def zeros(self) -> np.ndarray:
        return np.zeros((self.device_rows, self.CP), dtype=np.float32)

    def random(self, seed: int = 0, scale: float = 0.01) -> np.ndarray:
        rng = np.random.default_rng(seed)
        W = self.zeros()
        W[:, : self.R] = rng.standard_normal((self.device_rows, self.R)).astype(np.float32) * scale
        return W

    # -- dense math on the device layout (fp32/fp64 reference) -------------
    def logits(self, W: np.ndarray, X: np.ndarray) -> np.ndarray:
        W = np.asarray(W).reshape(self.device_rows, self.CP)
        return X @ W[: self.F, : self.R] + W[self.F, : self.R]

    def predict(self, W: np.ndarray, X: np.ndarray) -> np.ndarray:
        return np.argmax(self.logits(W, X), axis=1)
