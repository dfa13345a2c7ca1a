"""Sparse row containers shared by the producer, the ring buffer and the kernels.

``CSRBatch`` is the interchange format (indptr / indices / values / labels).
``pack_rows`` lays rows out the way the device ring stores them (see
``include/psb200.h``: ``PSB_ROW_*``): fixed-capacity ELL slots whose entries are
(int32 index, fp32 value) pairs grouped by *owner* = ``index % NOWN``; the per-row
group offsets let each warp of the gradient pass walk only the features it owns,
so the transposed product X^T R needs no atomics.

Reference for what a row is: ``LabeledData.java:14-28`` (sparse map + label) and
``CsvProducer.java:52-58`` (non-zero filter).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Tuple

import numpy as np

NOWN = 16               # gradient owners per row == warps of the worker CTA
HDR_WORDS = 12          # 48-byte slot header: nnz, label, id(lo,hi), 16 x u16 group start offsets


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


def slot_words(row_capacity: int) -> int:
    """32-bit words per ring slot: header + 2 words per entry."""
    return HDR_WORDS + 2 * row_capacity


def pack_rows(batch: CSRBatch, row_capacity: int, first_id: int = 1) -> np.ndarray:
    """Pack rows into device slot images: uint32[n, slot_words].

    Header words: [0]=nnz, [1]=label, [2]=insertion id low, [3]=insertion id high,
    [4..11] = 16 uint16 group start offsets (owner g covers entries off[g]..off[g+1], off[16]=nnz).
    Entry e: word 12+2e = int32 feature index, word 13+2e = fp32 value bits.
    """
    n = len(batch)
    sw = slot_words(row_capacity)
    out = np.zeros((n, sw), dtype=np.uint32)
    lens = np.diff(batch.indptr)
    if n and lens.max() > row_capacity:
        raise ValueError(f"row with {int(lens.max())} non-zeros exceeds the slot capacity {row_capacity}")
    for i in range(n):
        a, b = batch.indptr[i], batch.indptr[i + 1]
        idx = batch.indices[a:b]
        val = batch.values[a:b]
        own = idx % NOWN
        order = np.lexsort((idx, own))
        idx, val, own = idx[order], val[order], own[order]
        off = np.zeros(NOWN + 1, dtype=np.uint16)
        off[1:] = np.cumsum(np.bincount(own, minlength=NOWN))
        ins = first_id + i
        out[i, 0] = b - a
        out[i, 1] = np.uint32(batch.labels[i])
        out[i, 2] = ins & 0xFFFFFFFF
        out[i, 3] = (ins >> 32) & 0xFFFFFFFF
        # 16 group start offsets fit words 4..11; the 17th (== nnz) is word 0
        out[i, 4:12] = off[:NOWN].view(np.uint32)
        e = HDR_WORDS + 2 * np.arange(b - a)
        out[i, e] = idx.astype(np.int32).view(np.uint32)
        out[i, e + 1] = val.astype(np.float32).view(np.uint32)
    return out


def unpack_rows(slots: np.ndarray, num_features: int) -> CSRBatch:
    slots = np.asarray(slots, dtype=np.uint32)
    n = slots.shape[0]
    nnz = slots[:, 0].astype(np.int64)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(nnz, out=indptr[1:])
    indices = np.zeros(indptr[-1], dtype=np.int32)
    values = np.zeros(indptr[-1], dtype=np.float32)
    for i in range(n):
        e = HDR_WORDS + 2 * np.arange(nnz[i])
        idx = slots[i, e].view(np.int32)
        val = slots[i, e + 1].view(np.float32)
        order = np.argsort(idx, kind="stable")
        indices[indptr[i]:indptr[i + 1]] = idx[order]
        values[indptr[i]:indptr[i + 1]] = val[order]
    return CSRBatch(indptr, indices, values, slots[:, 1].astype(np.int32), num_features)


def insertion_ids(slots: np.ndarray) -> np.ndarray:
    slots = np.asarray(slots, dtype=np.uint32)
    return slots[:, 2].astype(np.int64) | (slots[:, 3].astype(np.int64) << 32)
