"""Metrics vs sklearn; CSV / slot packing round trips; synthetic data shape."""
import os

import numpy as np
from sklearn.metrics import accuracy_score, f1_score

from psb200.data import (CSRBatch, ReviewGenerator, TEST_SUPPORTS, header, pack_rows, random_csr, read_csv,
                         unpack_rows, write_csv)
from psb200.data.csr import insertion_ids
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


def test_slot_packing_groups_by_owner():
    b = random_csr(30, 256, 20, seed=3)
    slots = pack_rows(b, 32, first_id=7)
    assert np.array_equal(insertion_ids(slots), np.arange(7, 37))
    u = unpack_rows(slots, 256)
    assert np.allclose(u.to_dense(), b.to_dense()) and np.array_equal(u.labels, b.labels)
    for i in range(30):
        off = list(slots[i, 4:12].view(np.uint16)) + [slots[i, 0]]
        idx = slots[i, 12:12 + 2 * slots[i, 0]:2].view(np.int32)
        for g in range(16):
            assert np.all(idx[off[g]:off[g + 1]] % 16 == g)


def test_synthetic_shape():
    g = ReviewGenerator()
    te = g.test_split()
    assert len(te) == 4877 and tuple(np.bincount(te.labels)[1:]) == TEST_SUPPORTS
    d = te.slice(np.arange(50)).to_dense()
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-5)
    assert set(np.unique(te.labels)) == {1, 2, 3, 4, 5}
