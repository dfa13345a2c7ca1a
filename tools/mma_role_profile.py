"""MMA-warp time breakdown of the tcgen05 worker kernel (needs a -DPSB_TC_PROFILE build selected with
PSB200_LIB): cycles waiting for A chunks, for the residual tile, for B chunks, and the role total.
Usage: PSB200_LIB=build/exp_prof.so python tools/mma_role_profile.py [W ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

cap, F = 1024, 1024
dev = torch.device("cuda", 0)
for W in [int(a) for a in sys.argv[1:]] or [1, 148, 296]:
    X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
    eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap, local_solver=os.environ.get('SOLVER', 'lbfgs2')), W, impl=native.IMPL_TC, stage_rows=W)
    eng.fill(X, y)
    for _ in range(4):
        eng.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.step()
    e1.record()
    torch.cuda.synchronize()
    i = eng.info.cpu().float().mean(0).tolist()
    tot = i[7]
    print(f"W={W} kernel={e0.elapsed_time(e1) * 1e3:.0f} us  MMA warp (cycles/64, mean over workers): waitA={i[4]:.0f} "
          f"waitR={i[5]:.0f} waitB={i[6]:.0f} role_total={tot:.0f} issue+other={tot - i[4] - i[5] - i[6]:.0f}  evals={i[0]:.1f}")
    eng.close()
    del X, y
