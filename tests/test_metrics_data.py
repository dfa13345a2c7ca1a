"""Metrics vs sklearn; CSV / slot packing round trips; synthetic data shape."""
import os

import numpy as np
from sklearn.metrics import accuracy_score, f1_score

from psb200.data import (ReviewGenerator, TEST_SUPPORTS, header, pack_stage_rows, random_csr, read_csv,
                         stage_stride_words, unpack_stage_rows, write_csv)
from psb200.utils.metrics import evaluate


def test_weighted_f1_and_accuracy_match_sklearn():
    rng = np.random.default_rng(0)
    for _ in range(20):
        y = rng.integers(1, 6, size=500)
        p = np.where(rng.random(500) < 0.5, y, rng.integers(0, 6, size=500))
        m = evaluate(y, p, 6)
        assert abs(m.f1 - f1_score(y, p, average="weighted", labels=np.unique(y), zero_division=0)) < 1e-12
        assert abs(m.accuracy - accuracy_score(y, p)) < 1e-12


def test_f1_when_a_class_is_never_predicted():
    y = np.array([1, 1, 2, 2, 3])
    p = np.array([1, 1, 1, 1, 1])
    m = evaluate(y, p, 6)
    assert abs(m.f1 - f1_score(y, p, average="weighted", zero_division=0)) < 1e-12


def test_csv_roundtrip(tmp_path):
    b = ReviewGenerator(num_features=64, vocab=500, mean_len=12).train_split(40)
    path = os.path.join(tmp_path, "d.csv")
    write_csv(path, b)
    with open(path) as f:
        assert f.readline().strip() == header(64)
    r = read_csv(path)
    assert len(r) == 40 and np.array_equal(r.labels, b.labels)
    assert np.allclose(r.to_dense(), b.to_dense(), atol=1e-6)


def test_stage_packing_roundtrip():
    b = random_csr(30, 256, 20, seed=3)
    ws = np.stack([np.arange(30) % 4, np.arange(30) * 3 % 64], axis=1).astype(np.int32)
    stage = pack_stage_rows(b, ws, max_nnz=24)
    assert stage.shape == (30, stage_stride_words(24)) and np.all(stage[:, 0] == 20)
    u, ws2 = unpack_stage_rows(stage, 256)
    assert np.array_equal(ws2, ws) and np.array_equal(u.labels, b.labels)
    assert np.allclose(u.to_dense(), b.to_dense())


def test_synthetic_shape():
    g = ReviewGenerator()
    te = g.test_split()
    assert len(te) == 4877 and tuple(np.bincount(te.labels)[1:]) == TEST_SUPPORTS
    d = te.slice(np.arange(50)).to_dense()
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-5)
    assert set(np.unique(te.labels)) == {1, 2, 3, 4, 5}
