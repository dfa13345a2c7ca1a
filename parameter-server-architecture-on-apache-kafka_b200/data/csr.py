"""Sparse row containers shared by the producer, the engines and the kernels.

``CSRBatch`` is the interchange format (indptr / indices / values / labels).
``pack_stage_rows`` lays rows out the way the sparse ingestion kernel reads them (``psb_ring_scatter``,
``csrc/ps_ring.cu``): one fixed-stride record per row — nnz, label, worker, slot, then nnz (int32 index,
fp32 value) pairs — which the kernel expands into the dense fp16 ring slot.  For rows below ~25 % density
this is fewer PCIe bytes than shipping the dense fp16 row.

Reference for what a row is: ``LabeledData.java:14-28`` (sparse map + label) and
``CsvProducer.java:52-58`` (non-zero filter).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import numpy as np

STAGE_HDR_WORDS = 4     # nnz, label, worker, slot


@dataclass
class CSRBatch:
    indptr: np.ndarray      # int64[n+1]
    indices: np.ndarray     # int32[nnz]
    values: np.ndarray      # float32[nnz]
    labels: np.ndarray      # int32[n]
    num_features: int

    def __len__(self) -> int:
        return len(self.labels)

    @property
    def nnz(self) -> int:
        return int(self.indptr[-1])

    def row(self, i: int) -> Tuple[np.ndarray, np.ndarray, int]:
        a, b = self.indptr[i], self.indptr[i + 1]
        return self.indices[a:b], self.values[a:b], int(self.labels[i])

    def slice(self, rows) -> "CSRBatch":
        rows = np.asarray(rows, dtype=np.int64)
        lens = (self.indptr[rows + 1] - self.indptr[rows]).astype(np.int64)
        indptr = np.zeros(len(rows) + 1, dtype=np.int64)
        np.cumsum(lens, out=indptr[1:])
        if len(rows):
            gather = np.concatenate([np.arange(self.indptr[r], self.indptr[r + 1]) for r in rows]) \
                if indptr[-1] else np.zeros(0, dtype=np.int64)
        else:
            gather = np.zeros(0, dtype=np.int64)
        return CSRBatch(indptr, self.indices[gather], self.values[gather], self.labels[rows].copy(),
                        self.num_features)

    def to_dense(self, dtype=np.float32) -> np.ndarray:
        X = np.zeros((len(self), self.num_features), dtype=dtype)
        rows = np.repeat(np.arange(len(self)), np.diff(self.indptr))
        X[rows, self.indices] = self.values
        return X

    @staticmethod
    def from_dense(X: np.ndarray, labels: np.ndarray) -> "CSRBatch":
        X = np.asarray(X)
        mask = X != 0
        counts = mask.sum(axis=1)
        indptr = np.zeros(len(X) + 1, dtype=np.int64)
        np.cumsum(counts, out=indptr[1:])
        r, c = np.nonzero(mask)
        return CSRBatch(indptr, c.astype(np.int32), X[r, c].astype(np.float32),
                        np.asarray(labels, dtype=np.int32), X.shape[1])

    @staticmethod
    def concat(a: "CSRBatch", b: "CSRBatch") -> "CSRBatch":
        indptr = np.concatenate([a.indptr, b.indptr[1:] + a.indptr[-1]])
        return CSRBatch(indptr, np.concatenate([a.indices, b.indices]),
                        np.concatenate([a.values, b.values]),
                        np.concatenate([a.labels, b.labels]), a.num_features)


def stage_stride_words(max_nnz: int) -> int:
    """32-bit words per staged row: 4 header words + 2 per entry."""
    return STAGE_HDR_WORDS + 2 * max_nnz


def pack_stage_rows(batch: CSRBatch, worker_slot: np.ndarray, max_nnz: int = 0) -> np.ndarray:
    """Rows -> uint32[n, stride] records for ``psb_ring_scatter``; worker_slot is int[n, 2]."""
    n = len(batch)
    lens = np.diff(batch.indptr)
    max_nnz = max_nnz or (int(lens.max()) if n else 1)
    if n and lens.max() > max_nnz:
        raise ValueError(f"row with {int(lens.max())} non-zeros exceeds the staging capacity {max_nnz}")
    out = np.zeros((n, stage_stride_words(max_nnz)), dtype=np.uint32)
    out[:, 0] = lens
    out[:, 1] = batch.labels.astype(np.uint32)
    out[:, 2:4] = np.asarray(worker_slot, dtype=np.int32).view(np.uint32)
    rows = np.repeat(np.arange(n), lens)
    within = np.arange(batch.nnz) - np.repeat(batch.indptr[:-1], lens)
    out[rows, STAGE_HDR_WORDS + 2 * within] = batch.indices.astype(np.int32).view(np.uint32)
    out[rows, STAGE_HDR_WORDS + 2 * within + 1] = batch.values.astype(np.float32).view(np.uint32)
    return out


def unpack_stage_rows(stage: np.ndarray, num_features: int) -> Tuple[CSRBatch, np.ndarray]:
    """Inverse of pack_stage_rows: (batch, worker_slot)."""
    stage = np.asarray(stage, dtype=np.uint32)
    n = stage.shape[0]
    lens = stage[:, 0].astype(np.int64)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=indptr[1:])
    rows = np.repeat(np.arange(n), lens)
    within = np.arange(indptr[-1]) - np.repeat(indptr[:-1], lens)
    indices = stage[rows, STAGE_HDR_WORDS + 2 * within].view(np.int32)
    values = stage[rows, STAGE_HDR_WORDS + 2 * within + 1].view(np.float32)
    batch = CSRBatch(indptr, indices.copy(), values.copy(), stage[:, 1].astype(np.int32), num_features)
    return batch, stage[:, 2:4].view(np.int32).copy()
