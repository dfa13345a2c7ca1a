"""Fault injection on real GPUs: one rank is slowed on the host; how far do the others get?
sequential: everyone runs at the straggler's pace; bounded delay k: the fast ranks run k clocks ahead, then
the in-kernel gate holds them; eventual: unaffected.  torchrun launch; one JSON line on rank 0."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine, _view
from psb200.ops import native

p = argparse.ArgumentParser()
p.add_argument("--consistency_model", type=int, default=0)
p.add_argument("--steps", type=int, default=40)
p.add_argument("--delay_ms", type=float, default=5.0)
p.add_argument("--straggler", type=int, default=-1, help="rank to slow down (-1: last)")
p.add_argument("--workers_per_gpu", type=int, default=8)
a = p.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
strag = a.straggler if a.straggler >= 0 else world - 1
W, cap, F = a.workers_per_gpu, 256, 1024
cfg = PSConfig(num_workers=W * world, max_buffer_size=cap, consistency_model=a.consistency_model)
eng = DeviceEngine(cfg, W, rank, world, native.IMPL_TC, dev)
eng.connect()
X, y = synth_buffers(torch, W, cap, F, 0.45, 5 + rank, dev)
eng.fill(X, y)
clocks_view = _view(native.lib().psb_comm_rank_clocks(eng.block, F, world), world, dev, torch.int32)
dist.barrier()
torch.cuda.synchronize()
t0 = time.time()
max_gap = 0
for i in range(a.steps):
    if rank == strag:
        time.sleep(a.delay_ms / 1000.0)
    eng.step()
    torch.cuda.synchronize()
    if a.consistency_model != 0:
        c = clocks_view.cpu().numpy()
        max_gap = max(max_gap, int(c.max() - c.min()))
mine = time.time() - t0
status = eng.status()
times = [None] * world
gaps = [None] * world
dist.all_gather_object(times, mine)
dist.all_gather_object(gaps, max_gap)
if rank == 0:
    fast = [t for r, t in enumerate(times) if r != strag]
    print(json.dumps({"tool": "straggler_sweep", "consistency_model": a.consistency_model, "n_gpus": world,
                      "steps": a.steps, "straggler_rank": strag, "delay_ms_per_step": a.delay_ms,
                      "straggler_seconds": times[strag], "fast_ranks_seconds_mean": float(np.mean(fast)),
                      "fast_ranks_seconds_max": float(np.max(fast)), "max_observed_clock_gap": int(max(gaps)),
                      "status": status}))
dist.barrier()
eng.close()
dist.destroy_process_group()
