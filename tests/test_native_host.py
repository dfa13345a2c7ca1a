"""C++ host runtime vs its Python statement: buffer policy, tracker, CSV tokenizer."""
import os
import random

import numpy as np
import pytest

from psb200.consistency import ConsistencyScheduler, ProtocolViolation
from psb200.data import ReviewGenerator, read_csv, write_csv
from psb200.engine.buffer import DynamicBuffer
from psb200.ops import native


def test_native_buffer_matches_python_policy():
    rng = random.Random(4)
    py, cc = DynamicBuffer(16, 128, 0.3), native.NativeBuffer(16, 128, 0.3)
    t = 0.0
    for i in range(3000):
        if i % 400 == 0:
            gap = rng.choice([5.0, 50.0, 400.0, 2000.0])
        t += gap
        a, b = py.on_event(t), cc.on_event(t)
        assert a == b and (py.ring.head, py.ring.count, py.target) == (cc.head, cc.count, cc.target)
    assert cc.ring_total() == py.ring.total == 3000


def test_native_buffer_batch_form():
    bufs = [native.NativeBuffer(4, 32, 0.3) for _ in range(3)]
    ref = [DynamicBuffer(4, 32, 0.3) for _ in range(3)]
    workers = np.array([0, 1, 2, 0, 0, 1] * 20, dtype=np.int32)
    now = np.arange(len(workers)) * 700.0
    ws = np.zeros((len(workers), 2), dtype=np.int32)
    win = np.zeros((3, 2), dtype=np.int32)
    native.buffers_on_batch(bufs, workers, now, ws, win)
    later = {w: 0 for w in range(3)}                     # rows of the same worker that follow in the batch
    expect = [ref[w].on_event(now[i]) for i, w in enumerate(workers)]
    for i in range(len(workers) - 1, -1, -1):
        w = workers[i]
        assert ws[i, 0] == w and ws[i, 1] == (expect[i] if later[w] < 32 else -1)   # superseded rows are dropped
        later[w] += 1
    for w in range(3):
        assert (win[w, 0], win[w, 1]) == (ref[w].ring.head, ref[w].ring.count)


@pytest.mark.parametrize("order", ["grouped", "round_robin", "random"])
def test_batch_slot_assignment_equals_event_by_event_replay_and_reports_what_changed(order):
    """psb_buffers_on_batch2 groups the events by worker before it runs the policy; slots, windows, the per-worker
    change table (head / count before, rows added, count after) and the largest per-worker row count must be what an
    event-by-event replay gives, whatever the arrival order; rings with spare slots wrap at the ring size."""
    W, n_per = 5, 70
    rng = np.random.default_rng(3)
    workers = np.repeat(np.arange(W, dtype=np.int32), n_per)
    if order == "round_robin":
        workers = np.tile(np.arange(W, dtype=np.int32), n_per)
    elif order == "random":
        workers = rng.permutation(workers).astype(np.int32)
    now = np.cumsum(rng.choice([5.0, 40.0, 900.0], size=len(workers)))
    bufs = [native.NativeBuffer(8, 48, 0.3, ring=64) for _ in range(W)]
    ref = [native.NativeBuffer(8, 48, 0.3, ring=64) for _ in range(W)]
    for b, r in zip(bufs, ref):                       # some history first, so that "before" is not all zeros
        for t in (1.0, 2.0, 3.0):
            b.on_event(t), r.on_event(t)
    ws = np.zeros((len(workers), 2), dtype=np.int32)
    win = np.zeros((W, 2), dtype=np.int32)
    ch = np.zeros((W, 4), dtype=np.int32)
    most = native.buffers_on_batch(bufs, workers, now, ws, win, change=ch)
    before = [(r.head, r.count) for r in ref]
    slots = [ref[w].on_event(t) for w, t in zip(workers, now)]
    assert most == n_per and ws[:, 0].tolist() == workers.tolist()
    later = {w: 0 for w in range(W)}
    for i in range(len(workers) - 1, -1, -1):         # rows beyond the ring size are superseded (slot -1)
        w = int(workers[i])
        assert ws[i, 1] == (slots[i] if later[w] < 64 else -1)
        later[w] += 1
    for w in range(W):
        assert (win[w, 0], win[w, 1]) == (ref[w].head, ref[w].count)
        assert ch[w].tolist() == [before[w][0], before[w][1], n_per, ref[w].count]
        assert np.array_equal(bufs[w].get_state(), ref[w].get_state())


def test_batch_with_more_rows_than_the_ring_keeps_only_the_last_row_of_every_slot():
    """The store kernels write the rows of one batch in parallel, so a (worker, slot) pair may appear once."""
    cap = 8
    bufs = [native.NativeBuffer(4, cap, 0.3) for _ in range(2)]
    workers = np.array([0] * 21 + [1] * 5, dtype=np.int32)
    ws = np.zeros((len(workers), 2), dtype=np.int32)
    win = np.zeros((2, 2), dtype=np.int32)
    native.buffers_on_batch(bufs, workers, np.zeros(len(workers)), ws, win)
    live = ws[ws[:, 1] >= 0]
    assert len({(int(w), int(s)) for w, s in live}) == len(live)            # no slot written twice
    assert (ws[:21 - cap, 1] == -1).all() and (ws[21 - cap:, 1] >= 0).all()  # worker 0: only its last `cap` rows survive
    assert sorted(ws[21 - cap:21, 1].tolist()) == list(range(cap))
    assert (win[0, 0], win[0, 1]) == (21 % cap, cap) and (win[1, 0], win[1, 1]) == (5, 5)


@pytest.mark.parametrize("model", [0, 3, -1])
def test_native_tracker_matches_python_and_rejects_violations(model):
    t, s = native.NativeTracker(3, model), ConsistencyScheduler(3, model)
    assert sorted(t.on_gradient(1, 0)) == sorted(s.on_gradient(1, 0))
    with pytest.raises(native.NativeError, match="Expected value 1, actual value 0"):
        t.on_gradient(1, 0)
    with pytest.raises(ProtocolViolation):
        s.on_gradient(1, 0)
    assert t.clocks() == s.tracker.clocks


def test_csv_tokenizer_matches_numpy_reader(tmp_path):
    b = ReviewGenerator(num_features=128, vocab=800, mean_len=30).train_split(300)
    path = os.path.join(tmp_path, "d.csv")
    write_csv(path, b)
    a, r = native.csv_to_csr(path), read_csv(path)
    assert np.array_equal(a.indptr, r.indptr) and np.array_equal(a.indices, r.indices)
    assert np.allclose(a.values, r.values, atol=1e-7) and np.array_equal(a.labels, r.labels)
    with open(path, "a") as f:                       # exponent notation and a CRLF line go through strtof
        f.write(",".join(["1e-3"] + ["0"] * 126 + ["-2.5E-1"]) + ",4\r\n")
    a = native.csv_to_csr(path)
    assert len(a) == 301 and np.allclose(a.row(300)[1], [1e-3, -0.25]) and a.row(300)[2] == 4


def test_one_pass_loader_equals_the_two_pass_loader_for_every_thread_count(tmp_path):
    """psb_csv_open / shape / fill / close: the file mapped once and tokenised by several threads, each on its own
    chunk cut at a line end — the same CSR as psb_csv_count + psb_csv_parse, whatever the number of threads; files
    that are empty, hold only a header, or lack the final newline."""
    b = ReviewGenerator(num_features=1024, mean_len=60).train_split(2200)          # ~5 MB of text: up to 4 chunks
    path = os.path.join(tmp_path, "big.csv")
    write_csv(path, b)
    with open(path, "rb+") as f:                                                   # no newline at the end
        f.seek(-1, os.SEEK_END)
        f.truncate()
    ref = native.csv_to_csr_two_pass(path)
    assert len(ref) == 2200 and ref.num_features == 1024
    for threads in (1, 2, 3, 4, 0):
        a = native.csv_to_csr(path, threads=threads)
        assert np.array_equal(a.indptr, ref.indptr) and np.array_equal(a.indices, ref.indices), threads
        assert np.array_equal(a.values, ref.values) and np.array_equal(a.labels, ref.labels) and a.num_features == 1024
    empty, head = os.path.join(tmp_path, "empty.csv"), os.path.join(tmp_path, "head.csv")
    open(empty, "w").close()
    open(head, "w").write("0,1,2,Score\n")
    for path2 in (empty, head):
        a = native.csv_to_csr(path2)
        assert len(a) == 0 and a.indptr.tolist() == [0]
    with pytest.raises(native.NativeError):
        native.csv_to_csr(os.path.join(tmp_path, "nope.csv"))


def test_csv_tokenizer_field_forms(tmp_path):
    """Every way a field is taken: the 8-byte run skip of "0,0,0,0," at every alignment, "0.0", signs, blanks,
    leading zeros, long digit strings (strtof beyond 18 digits), exponents, a zero label and a missing final newline
    (CsvProducer.java:46-58 splits on ',' and parses floats).  Values must be the correctly rounded float32 of the
    text wherever the mantissa fits a double (<= 15 digits), and within 1 ulp beyond."""
    rng = np.random.default_rng(2)
    forms = ["0", "0.0", "-0", "+0.50", "1", "-1", "007.25", " 3.5", "2.5 ", ".5", "5.", "-.125", "1e-3", "-2.5E+1",
             "0.000021181", "123456789012345678", "1234567890123456789012", "0.1234567890123456789", "0e5", "",
             # the eight-digits-at-a-time path and its exits: 0 / 7 / 8 / 9 / 15 / 16 fraction digits, no ',' behind
             "1.", "0.5", "-0.25", "9.99", "0.1234567", "0.12345678", "-0.123456789", "0.123456789012345",
             "0.1234567890123456", "3.00000001", "0.5 ", "-0.021181"]
    lines, want = [], []
    for r in range(200):
        fields = ["0"] * 40
        for c in rng.choice(40, size=rng.integers(0, 9), replace=False):
            fields[c] = forms[rng.integers(0, len(forms))]
        if r == 7:
            fields = ["0"] * 40                                   # an all-zero row
        label = str(int(rng.integers(0, 6)))
        lines.append(",".join(fields + [label]))
        want.append(([float(f.strip() or 0) for f in fields], int(label)))
    path = os.path.join(tmp_path, "forms.csv")
    with open(path, "w") as f:
        f.write(",".join(str(i) for i in range(40)) + ",Score\n" + "\n".join(lines))     # no newline at the end
    a = native.csv_to_csr(path)
    assert len(a) == 200 and a.num_features == 40
    for r, (vals, label) in enumerate(want):
        idx, v, lab = a.row(r)
        ref = np.asarray(vals, dtype=np.float32)
        assert lab == label and idx.tolist() == np.nonzero(ref)[0].tolist()
        assert np.allclose(v, ref[ref != 0], rtol=1.2e-7, atol=0)


def test_bitmap_packers_roundtrip_dense_and_csr():
    """psb_pack_bitmap_dense / _csr: presence bitmap + fp16 values reproduce the fp16 rows bit for bit
    (sparse wire format of ingestion; LabeledData.java:14-28)."""
    from psb200.data.synthetic import random_csr
    rng = np.random.default_rng(5)
    F, n = 256, 37
    dense = (rng.standard_normal((n, F)) * (rng.random((n, F)) < 0.3)).astype(np.float16)
    dense[3] = 0                                    # an empty row
    dense[4, 7] = np.float16(-0.0)                  # negative zero is "absent"
    labels = rng.integers(1, 6, n).astype(np.int32)
    pk = native.PackedRows(64, F, pin=False).pack_dense(dense, labels)
    assert pk.n == n and pk.nnz == int((dense != 0).sum())
    assert np.array_equal(pk.to_dense().view(np.uint16), np.where(dense == 0, np.float16(0), dense).view(np.uint16))
    assert pk.row_off[:n + 1].tolist() == np.concatenate([[0], np.cumsum((dense != 0).sum(1))]).tolist()
    assert pk.labels[:n].tolist() == labels.tolist()
    assert pk.nbytes == n * F // 8 + (n + 1) * 4 + pk.nnz * 2 + n * 4

    b = random_csr(29, F, 20, seed=3)
    perm = b.indices.copy()                          # unsorted columns inside a row must not matter
    for r in range(len(b)):
        a0, a1 = b.indptr[r], b.indptr[r + 1]
        o = rng.permutation(a1 - a0)
        perm[a0:a1] = b.indices[a0:a1][o]
        b.values[a0:a1] = b.values[a0:a1][o]
    b.indices[:] = perm
    pk2 = native.PackedRows(32, F, pin=False).pack_csr(b)
    assert np.array_equal(pk2.to_dense(), b.to_dense(np.float16))


def test_palette_packer_roundtrip_with_raw_fallback():
    """psb_pack_palette_dense: per-row palette + 4-bit codes reproduce hashed-count rows bit for bit; a row with
    more than 16 distinct values, an empty row and an odd number of non-zeros take the fallback / padding paths."""
    rng = np.random.default_rng(11)
    F, n = 256, 45
    cnt = rng.integers(1, 4, (n, F)) * (rng.integers(0, 2, (n, F)) * 2 - 1) * (rng.random((n, F)) < 0.4)
    rows = (cnt / np.maximum(np.linalg.norm(cnt, axis=1, keepdims=True), 1e-6)).astype(np.float16)
    rows[5] = 0
    rows[7] = (rng.standard_normal(F) * (rng.random(F) < 0.5)).astype(np.float16)        # > 16 distinct values: raw
    rows[9] = 0
    rows[9, [3, 77, 200]] = np.float16(0.5)                                               # 3 non-zeros, 1 palette entry
    labels = rng.integers(1, 6, n).astype(np.int32)
    pk = native.PackedRows(64, F, pin=False, palette=True).pack_dense(rows, labels)
    ref = np.where(rows == 0, np.float16(0), rows)
    assert np.array_equal(pk.to_dense().view(np.uint16), ref.view(np.uint16))
    off = pk.row_off[:n + 1].numpy()
    assert (off % 2 == 0).all() and off[n] == pk.payload_bytes
    assert int(pk.payload[off[7]:off[7] + 2].numpy().view(np.uint16)[0]) == 0            # raw row
    assert int(pk.payload[off[9]:off[9] + 2].numpy().view(np.uint16)[0]) == 1
    plain = native.PackedRows(64, F, pin=False).pack_dense(rows, labels)
    assert pk.nbytes < 0.6 * plain.nbytes
