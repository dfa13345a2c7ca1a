"""ncu target covering every hot kernel once: tcgen05 worker step (296 workers), SIMT worker step, tcgen05 and
SIMT evaluation, ring ingestion, sharded sparse step + eval.  Usage: python tools/profile_all.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.engine.sparse_engine import SparseShardedEngine
from psb200.ops import native
from tools.bench_sparse import Batch

dev = torch.device("cuda", 0)
cap, F = 1024, 1024
for W, impl in ((296, native.IMPL_TC), (148, native.IMPL_SIMT)):
    X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
    eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap), W, impl=impl, stage_rows=W * 128)
    eng.fill(X, y)
    eng.test_X, eng.test_y = X[:5].reshape(-1, F)[:4877].contiguous(), y[:5].reshape(-1)[:4877].contiguous()
    rows = X[0, :128].repeat(W, 1).cpu().pin_memory()
    labs = y[0, :128].repeat(W).cpu().pin_memory()
    workers = np.repeat(np.arange(W, dtype=np.int32), 128)
    packs = [native.PackedRows(W * 128, F, palette=pal).pack_dense(rows, labs) for pal in (True, False)]
    for it in range(3):
        eng.ingest(rows if it == 0 else packs[it - 1], labs, workers, np.zeros(W * 128))   # dense, palette, bitmap
        eng.step()
        eng.evaluate()
    print("dense impl", impl, "W", W, "loss", float(eng.read_loss().mean()), "status", eng.status())
    eng.close()
    del X, y
# statistics kept at ingestion time, all workers' local models scored in one launch, sparse-row kernel + ELL store
eng = DeviceEngine(PSConfig(num_workers=148, max_buffer_size=cap), 148, impl=native.IMPL_TC, stage_rows=148 * 128,
                   ring_spare_rows=128)
X, y = synth_buffers(torch, 148, cap, F, 0.45, 1, dev)
eng.fill(X, y)
eng.test_X, eng.test_y = X[:5].reshape(-1, F)[:4877].contiguous(), y[:5].reshape(-1)[:4877].contiguous()
rows = X[0, :128].repeat(148, 1).cpu().pin_memory()
labs = y[0, :128].repeat(148).cpu().pin_memory()
pk = native.PackedRows(148 * 128, F, palette=True).pack_dense(rows, labs)
for it in range(2):
    eng.ingest(pk, None, np.repeat(np.arange(148, dtype=np.int32), 128), np.zeros(148 * 128))
    eng.step()
    eng.evaluate_workers()
print("stats-at-ingest loss", float(eng.read_loss().mean()), "status", eng.status())
eng.close()
from psb200.data import CSRBatch
from psb200.engine.sparse_row_engine import SparseRowEngine
sr = SparseRowEngine(PSConfig(num_workers=148, num_features=4096, max_buffer_size=256), 148, 320)
rng = np.random.default_rng(0)
d = ((rng.random((148 * 200, 4096)) < 0.05) * rng.integers(1, 4, size=(148 * 200, 4096))).astype(np.float32)
d /= np.maximum(np.linalg.norm(d, axis=1, keepdims=True), 1e-6)
sr.ingest(CSRBatch.from_dense(d, rng.integers(1, 6, size=148 * 200).astype(np.int32)), np.repeat(np.arange(148), 200))
for _ in range(2):
    sr.step()
print("sparse rows loss", float(sr.read_loss().mean()), "status", sr.status())
sr.close()
del X, y, d
sp = SparseShardedEngine(1 << 24, 5, 0, 1, 1, 0.5, -1, dev)
b = Batch(16384, 1 << 24, 128, 3, dev)
for _ in range(3):
    sp.step(b)
print("sparse loss", sp.mean_loss(b), "f1/acc", sp.evaluate(b))
sp.close()
