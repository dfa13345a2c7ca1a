"""Device times of the ingestion kernels at the headline shape (296 workers x 128 fresh rows, F = 1024, density 0.45):
ring store per wire format and the window-statistics update, CUDA events around 20 back-to-back launches each.
Usage: python tools/ingest_times.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

W, r, F, cap = 296, 128, 1024, 1024
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
rows = (torch.rand((W * r, F), generator=g) < 0.45).float() * torch.randint(1, 4, (W * r, F), generator=g).float() \
    * (torch.randint(0, 2, (W * r, F), generator=g).float() * 2 - 1)
rows = (rows / rows.norm(dim=1, keepdim=True).clamp_min(1e-6)).half().pin_memory()
labs = torch.randint(1, 6, (W * r,), generator=g, dtype=torch.int32).pin_memory()
workers = np.repeat(np.arange(W, dtype=np.int32), r)
eng = DeviceEngine(PSConfig(num_workers=W, max_buffer_size=cap), W, impl=native.IMPL_TC, stage_rows=W * r,
                   ring_spare_rows=r)
out = {}
for name, batch in (("dense", rows), ("bitmap", native.PackedRows(W * r, F).pack_dense(rows, labs)),
                    ("palette", native.PackedRows(W * r, F, palette=True).pack_dense(rows, labs))):
    for _ in range(3):
        eng.ingest(batch, labs, workers)
    torch.cuda.synchronize()
    j = (eng.stage_head - 1) % eng.stage_sets                    # the set that was just stored: replay its kernels
    s = torch.cuda.current_stream()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    for _ in range(20):
        if name == "palette":
            native.ring_store_palette(eng.stage_bitmap_dev[j], eng.stage_off_dev[j], eng.stage_dev[j], eng.stage_lab_dev[j],
                                      eng.stage_ws_dev[j], W * r, eng.X, eng.labels, eng.cap, F)
        elif name == "bitmap":
            native.ring_store_bitmap(eng.stage_bitmap_dev[j], eng.stage_off_dev[j], eng.stage_dev[j], eng.stage_lab_dev[j],
                                     eng.stage_ws_dev[j], W * r, eng.X, eng.labels, eng.cap, F)
        else:
            native.ring_store_dense(eng.stage_dev[j], eng.stage_lab_dev[j], eng.stage_ws_dev[j], W * r, eng.X, eng.labels,
                                    eng.cap, F)
    e1.record()
    for _ in range(20):
        native.window_stats_update(eng.X, eng.stage_change_dev[j], W, eng.cap, F, eng.stat_sums, eng._inv[0])
    e2.record()
    torch.cuda.synchronize()
    out[name] = {"store_us": e0.elapsed_time(e1) * 50.0, "stats_update_us": e1.elapsed_time(e2) * 50.0}
print(json.dumps({"tool": "ingest_times", "workers": W, "rows_per_worker": r, "features": F, **out}))
eng.close()
