"""ncu target: the headline shape only (296 workers x 1024 rows x 1024 features, tcgen05 worker step, sequential).
Usage: python tools/profile_worker.py [steps] [workers]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
W = int(sys.argv[2]) if len(sys.argv) > 2 else 296
cap, F = 1024, 1024
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap), W, impl=native.IMPL_TC)
eng.fill(X, y)
for _ in range(steps):
    eng.step()
print("loss", float(eng.read_loss().mean()), "status", eng.status())
eng.close()
