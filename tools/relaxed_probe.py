"""Per-launch device times of the per-worker relaxed modes (torchrun, one rank per GPU): eventual consistency,
W workers per GPU, `--cpl` clocks per launch.  Prints each rank's launch times (ms) and the in-kernel cycle counters."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from bench import synth_buffers
from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.ops import native

p = argparse.ArgumentParser()
p.add_argument("--cpl", type=int, default=8)
p.add_argument("--workers", type=int, default=296)
p.add_argument("--launches", type=int, default=6)
p.add_argument("--model", type=int, default=-1)
p.add_argument("--delay_worker", type=int, default=-1)
p.add_argument("--delay_cycles", type=int, default=0)
p.add_argument("--check", action="store_true", help="synchronise after every launch and report failed barrier waits at once")
p.add_argument("--spare", type=int, default=0, help="spare ring rows per worker (as bench.py's end-to-end engine)")
a = p.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
W, cap, F = a.workers, 1024, 1024
cfg = PSConfig(num_workers=W * world, num_features=F, consistency_model=a.model, max_buffer_size=cap)
eng = DeviceEngine(cfg, W, rank, world, native.IMPL_TC, dev, async_granularity="worker", max_clocks_per_launch=a.cpl,
                   ring_spare_rows=a.spare, stats_at_ingest=False, stage_rows=W * 128)
eng.connect()
X, y = synth_buffers(torch, W, cap, F, 0.45, 1234 + rank, dev)
eng.fill(X, y)
del X, y
dist.barrier()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.launches + 1)]
if a.delay_cycles:
    eng.inject_delay(a.delay_worker, a.delay_cycles)
ev[0].record()
for i in range(a.launches):
    eng.step(num_clocks=a.cpl)
    ev[i + 1].record()
    if a.check:
        torch.cuda.synchronize()
        inf = eng.info.cpu().numpy()
        for w in np.nonzero(inf[:, 2] == -2)[0]:
            print(f"rank {rank}: launch {i} worker {w} FAILED barrier wait site {-inf[w, 1]} evals {inf[w, 0]} "
                  f"launch ms {ev[i].elapsed_time(ev[i + 1]):.1f}", flush=True)
torch.cuda.synchronize()
ms = [round(ev[i].elapsed_time(ev[i + 1]), 2) for i in range(a.launches)]
info = eng.info.cpu().numpy()
h = eng.read_hist()
dist.barrier()
slow = [(i, t) for i, t in enumerate(ms) if t > 3 * float(np.median(ms))]
bad = np.nonzero(info[:, 2] == -2)[0]
failed = [(int(w), "site", int(-info[w, 1]), "evals", int(info[w, 0]), "clocks done", int(eng.worker_clocks()[rank * W + w]))
          for w in bad[:6]]
if len(ms) > 30:
    ms = ms[:3] + ["..."] + [f"median {float(np.median([t for t in ms if not isinstance(t, str)])):.2f}"]
print(f"rank {rank}: slow launches {slow} workers with a failed barrier wait {failed} launches ms {ms}  last-clock cycles/64 stats,solve,push,total {info[:, 4:8].mean(axis=0).round(0).tolist()} "
      f"lag max {int(h['lag'].max())} status {eng.status()}", flush=True)
eng.close()
dist.destroy_process_group()
