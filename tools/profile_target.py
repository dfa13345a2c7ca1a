"""Small fixed workload for ncu / launch listings: 148 workers x 1024 rows tcgen05 step, SIMT step,
tcgen05 + SIMT evaluation, ring ingestion.  Usage: python tools/profile_target.py [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
W, cap, F = 148, 1024, 1024
cfg = PSConfig(num_workers=W, max_buffer_size=cap)
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
for impl in (native.IMPL_TC, native.IMPL_SIMT):
    eng = DeviceEngine(cfg, W, impl=impl, stage_rows=W * 128)
    eng.fill(X, y)
    eng.test_X, eng.test_y = X[:5].reshape(-1, F)[:4877].contiguous(), y[:5].reshape(-1)[:4877].contiguous()
    rows = X[0, :128].repeat(W, 1).cpu().pin_memory()
    labs = y[0, :128].repeat(W).cpu().pin_memory()
    workers = np.repeat(np.arange(W, dtype=np.int32), 128)
    for _ in range(steps):
        eng.ingest(rows, labs, workers, np.zeros(W * 128))
        eng.step()
        eng.evaluate()
    print("impl", impl, "loss", float(eng.read_loss().mean()), "status", eng.status())
    eng.close()
