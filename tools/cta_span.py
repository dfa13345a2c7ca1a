"""Per-CTA start / end wall-clock stamps of one worker-kernel launch (needs a -DPSB_TC_TRACE build, PSB200_LIB)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

W = int(sys.argv[1]) if len(sys.argv) > 1 else 296
dev = torch.device("cuda", 0)
X, y = synth_buffers(torch, W, 1024, 1024, 0.45, 1, dev)
eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=1024), W, impl=native.IMPL_TC, stage_rows=W)
eng.fill(X, y)
for _ in range(3):
    eng.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
eng.step()
e1.record()
torch.cuda.synchronize()
i = eng.info.cpu().numpy().astype(np.int64)
st, en = i[:, 4], i[:, 5]
t0 = st.min()
print(f"W={W} event time {e0.elapsed_time(e1) * 1e3:.0f} us; CTA starts {(st - t0).min() / 1e3:.1f}..{(st - t0).max() / 1e3:.1f} us, "
      f"ends {(en - t0).min() / 1e3:.1f}..{(en - t0).max() / 1e3:.1f} us, mean duration {(en - st).mean() / 1e3:.1f} us")
order = np.argsort(st)
print("  start offsets (us) of blocks 0,1,147,148,149,295:", [round(float(st[b] - t0) / 1e3, 1) for b in (0, 1, 147, 148, 149, W - 1) if b < W])
eng.close()
