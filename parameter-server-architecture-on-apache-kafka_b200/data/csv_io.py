"""Dataset CSV reader / writer (pure NumPy fallback; the hot parser is native).

Format (``LogisticRegressionTaskSpark.java:77-90``, ``CsvProducer.java:41-58``):
header ``0,1,...,F-1,Score``; F float columns; last column integer label.
The reader tolerates a missing header when ``has_header=False`` (the reference
producer always skips one line).  ``psb200.ops.native.csv_to_csr`` is the C++
tokenizer used by the producer thread; this module is its oracle and the writer.
"""
from __future__ import annotations

import io
from typing import Iterator, Tuple

import numpy as np

from .csr import CSRBatch


def header(num_features: int) -> str:
    return ",".join(str(i) for i in range(num_features)) + ",Score"


def _fmt(v: float) -> str:
    if v == 0:
        return "0"
    return repr(round(float(v), 6))


def write_csv(path: str, batch: CSRBatch) -> None:
    F = batch.num_features
    with open(path, "w") as f:
        f.write(header(F) + "\n")
        for i in range(len(batch)):
            idx, val, lab = batch.row(i)
            cols = ["0"] * F
            for j, v in zip(idx, val):
                cols[j] = _fmt(v)
            f.write(",".join(cols) + "," + str(lab) + "\n")


def parse_line(line: str) -> Tuple[np.ndarray, np.ndarray, int]:
    """One CSV row -> (indices, values, label); zeros dropped (CsvProducer.java:52-57)."""
    parts = line.rstrip("\r\n").split(",")
    vals = np.asarray(parts[:-1], dtype=np.float32)
    nz = np.nonzero(vals)[0]
    return nz.astype(np.int32), vals[nz], int(float(parts[-1]))


def iter_rows(path: str, has_header: bool = True) -> Iterator[Tuple[np.ndarray, np.ndarray, int]]:
    with open(path, "r") as f:
        if has_header:
            f.readline()
        for line in f:
            if line.strip():
                yield parse_line(line)


def read_csv(path: str, has_header: bool = True, num_features: int = 0) -> CSRBatch:
    with open(path, "r") as f:
        first = f.readline() if has_header else None
        body = f.read()
    if not body.strip():
        F = num_features or (first.count(",") if first else 0)
        return CSRBatch(np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float32),
                        np.zeros(0, np.int32), F)
    arr = np.loadtxt(io.StringIO(body), delimiter=",", dtype=np.float64, ndmin=2)
    X = arr[:, :-1].astype(np.float32)
    y = arr[:, -1].astype(np.int32)
    return CSRBatch.from_dense(X, y)
