"""ncu target: N steps of the tcgen05 worker kernel at W workers per GPU.  Usage: profile_step.py W [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

W = int(sys.argv[1]) if len(sys.argv) > 1 else 148
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
cap, F = 1024, 1024
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap), W, impl=native.IMPL_TC, stage_rows=W)
eng.fill(X, y)
for _ in range(steps):
    eng.step()
print("W", W, "loss", float(eng.read_loss().mean()), "status", eng.status())
eng.close()
