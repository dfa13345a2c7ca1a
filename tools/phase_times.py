"""In-kernel phase times of the tcgen05 worker kernel (info[4..7], cycles/64) with and without the fused
standardisation statistics.  Usage: phase_times.py [W] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

W = int(sys.argv[1]) if len(sys.argv) > 1 else 296
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cap, F = 1024, 1024
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
for std in ((True,) if os.environ.get('ONLY_STD') else (True, False)):
    eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap, standardize=std), W, impl=native.IMPL_TC,
                       stage_rows=W)
    eng.fill(X, y)
    for _ in range(5):
        eng.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.step()
    e1.record()
    torch.cuda.synchronize()
    tot = eng.info.cpu().float()[:, 7]
    print(f"  per-CTA total cycles/64: min {tot.min():.0f} mean {tot.mean():.0f} max {tot.max():.0f}")
    info = eng.info.cpu().float().mean(0).tolist()
    print(f"standardize={std} W={W} ms/clock={e0.elapsed_time(e1) / steps:.4f} evals={info[0]:.2f} "
          f"cycles64 first={info[4]:.0f} solve={info[5]:.0f} push={info[6]:.0f} total={info[7]:.0f} status={eng.status()}")
    eng.close()
