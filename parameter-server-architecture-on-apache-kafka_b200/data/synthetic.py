"""Synthetic stand-in for the fine-food-reviews embedding CSVs.

The real files are not shipped (``.MISSING_LARGE_BLOBS:1-2``).  What is known of
their shape: header ``0,1,...,1023,Score`` (``LogisticRegressionTaskSpark.java:77-90``),
1024 signed hashed bag-of-words counts per row, L2-normalised (values are integer
ratios such as 0.021181, ``evaluation/python-ground-truth-algorithm.ipynb`` cell 6),
integer label 1..5 (``CsvProducer.java:58``); test split of 4,877 rows with supports
1038/741/1041/1024/1033 (``README.md:223-233``); a linear model reaches weighted
F1 ~0.47 on it.

Generator: documents are token draws from a Zipf vocabulary tilted per class;
tokens are hashed to (bucket, sign) the way a HashingVectorizer does; counts are
L2-normalised.  ``class_signal`` tunes how separable the classes are.
"""
from __future__ import annotations

import numpy as np

from .csr import CSRBatch

TEST_SUPPORTS = (1038, 741, 1041, 1024, 1033)       # README.md:223-233, labels 1..5


class ReviewGenerator:
    def __init__(self, num_features: int = 1024, num_classes: int = 5, vocab: int = 20000,
                 mean_len: float = 60.0, class_signal: float = 0.24, seed: int = 0):
        self.F, self.C, self.V = num_features, num_classes, vocab
        self.mean_len = mean_len
        rng = np.random.default_rng(seed + 7919)
        base = 1.0 / np.arange(1, vocab + 1) ** 1.05
        # ordinal structure: neighbouring star ratings share vocabulary, like real reviews
        latent = rng.standard_normal((vocab, 2)).astype(np.float64)
        angle = np.linspace(-1.0, 1.0, num_classes)
        tilt = np.stack([latent[:, 0] * a + latent[:, 1] * (1 - abs(a)) for a in angle], axis=0)
        logits = np.log(base)[None, :] + class_signal * tilt
        p = np.exp(logits - logits.max(axis=1, keepdims=True))
        self.cdf = np.cumsum(p / p.sum(axis=1, keepdims=True), axis=1)
        h = rng.integers(0, 2 ** 31, size=vocab)
        self.bucket = (h % num_features).astype(np.int32)
        self.sign = np.where((h >> 20) & 1, 1.0, -1.0).astype(np.float32)

    def sample(self, labels: np.ndarray, seed: int = 0) -> CSRBatch:
        """labels: int array with values 1..num_classes."""
        rng = np.random.default_rng(seed)
        labels = np.asarray(labels, dtype=np.int32)
        n = len(labels)
        lens = np.maximum(3, rng.lognormal(np.log(self.mean_len) - 0.32, 0.8, size=n)).astype(np.int64)
        lens = np.minimum(lens, 2000)
        indptr = [0]
        idx_chunks, val_chunks = [], []
        for i in range(n):
            u = rng.random(lens[i])
            tok = np.searchsorted(self.cdf[labels[i] - 1], u)
            tok = np.minimum(tok, self.V - 1)
            dense = np.bincount(self.bucket[tok], weights=self.sign[tok], minlength=self.F)
            nz = np.nonzero(dense)[0]
            v = dense[nz]
            nrm = np.sqrt((v * v).sum())
            if nrm == 0:                       # all counts cancelled: keep one token so the row is not empty
                nz = np.array([self.bucket[tok[0]]])
                v = np.array([1.0])
                nrm = 1.0
            idx_chunks.append(nz.astype(np.int32))
            val_chunks.append((v / nrm).astype(np.float32))
            indptr.append(indptr[-1] + len(nz))
        return CSRBatch(np.asarray(indptr, dtype=np.int64), np.concatenate(idx_chunks),
                        np.concatenate(val_chunks), labels.copy(), self.F)

    def train_split(self, n: int, seed: int = 1) -> CSRBatch:
        """Balanced, shuffled stream (the reference caps each label at 20,000 rows, README.md:215)."""
        rng = np.random.default_rng(seed)
        labels = (np.arange(n) % self.C) + 1
        rng.shuffle(labels)
        return self.sample(labels, seed=seed + 1)

    def test_split(self, n: int = 4877, seed: int = 2) -> CSRBatch:
        rng = np.random.default_rng(seed)
        if n == sum(TEST_SUPPORTS) and self.C == 5:
            labels = np.concatenate([np.full(s, k + 1) for k, s in enumerate(TEST_SUPPORTS)])
        else:
            labels = (np.arange(n) % self.C) + 1
        rng.shuffle(labels)
        return self.sample(labels, seed=seed + 1)


def random_csr(n: int, num_features: int, nnz_per_row: int, num_classes: int = 5, seed: int = 0) -> CSRBatch:
    """Fast structureless sparse batch for throughput runs and kernel tests
    (uniform feature ids, Gaussian values, L2-normalised rows, labels 1..C)."""
    rng = np.random.default_rng(seed)
    k = min(nnz_per_row, num_features)
    if num_features <= 4 * k:
        idx = np.stack([np.sort(rng.choice(num_features, size=k, replace=False)) for _ in range(n)])
    else:
        idx = np.sort(rng.integers(0, num_features, size=(n, k)), axis=1)
        dup = np.concatenate([np.zeros((n, 1), bool), idx[:, 1:] == idx[:, :-1]], axis=1)
        while dup.any():                      # re-draw duplicates until every row has distinct ids
            idx[dup] = rng.integers(0, num_features, size=int(dup.sum()))
            idx = np.sort(idx, axis=1)
            dup = np.concatenate([np.zeros((n, 1), bool), idx[:, 1:] == idx[:, :-1]], axis=1)
    val = rng.standard_normal((n, k)).astype(np.float32)
    val /= np.linalg.norm(val, axis=1, keepdims=True)
    indptr = np.arange(n + 1, dtype=np.int64) * k
    labels = rng.integers(1, num_classes + 1, size=n).astype(np.int32)
    return CSRBatch(indptr, idx.reshape(-1).astype(np.int32), val.reshape(-1), labels, num_features)
