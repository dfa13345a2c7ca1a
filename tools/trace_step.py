"""Timeline of block 0 of the tcgen05 worker kernel (needs a -DPSB_TC_TRACE build selected with PSB200_LIB).
Usage: PSB200_LIB=build/exp_trace.so python tools/trace_step.py [W]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

W = int(sys.argv[1]) if len(sys.argv) > 1 else 1
cap, F = 1024, 1024
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, cap, F, 0.45, 1, dev)
eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap), W, impl=native.IMPL_TC, stage_rows=W)
eng.fill(X, y)
for _ in range(3):
    eng.step()
torch.cuda.synchronize()
print(f"---- W={W}: ids 0 start, 1 tmem/barriers done, 2 pulled, 3 first eval done, 4 scaled, 10/11 direction setup, "
      "12/13 trial eval, 14 dot, 15 pair update, 5 solve done, 6 delta, 7 pushed; 20/21/22 pass staging/roles/tail", flush=True)
eng.close()
