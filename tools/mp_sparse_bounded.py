"""Bounded delay / eventual consistency on the sharded sparse path, real GPUs (torchrun, one rank per GPU).

The last rank is a straggler (sleeps --delay_ms on the host before every step).  Bounded delay k: the other ranks'
kernels wait inside the gate `min over ranks of completed steps >= v - k`, so after their k free steps they advance at
the straggler's pace; eventual: they finish long before it.  Prints MPSPARSE-BOUNDED ... OK|FAIL on rank 0."""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from psb200.data import random_csr
from psb200.engine.sparse_engine import DeviceCSR, SparseShardedEngine

p = argparse.ArgumentParser()
p.add_argument("--model", type=int, default=2)
p.add_argument("--steps", type=int, default=12)
p.add_argument("--delay_ms", type=float, default=25.0)
a = p.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
F, B = 60000, 512
eng = SparseShardedEngine(F, 5, rank, world, world, 0.5, a.model, dev)
eng.connect()
straggler = world - 1
if rank == straggler:
    eng.host_delay_s = a.delay_ms / 1000.0
batches = [DeviceCSR(random_csr(B, F, 40, seed=10 * s + rank), dev) for s in range(a.steps)]
dist.barrier()
t0 = time.time()
lags, losses = [], []
for v in range(a.steps):
    eng.step(batches[v])
    torch.cuda.synchronize()
    losses.append(eng.mean_loss(batches[v]))
    done = eng.completed_steps()
    lags.append(v - int(done.min()) if a.model > 0 else 0)       # (the gate saw an even smaller lag than this)
elapsed = time.time() - t0
dist.barrier()
out = [None] * world
dist.all_gather_object(out, {"elapsed": elapsed, "lag": max(lags), "status": eng.status(),
                             "finite": bool(np.isfinite(losses).all()), "first": losses[0], "last": losses[-1]})
ok = True
if rank == 0:
    total_delay = a.steps * a.delay_ms / 1000.0
    fast = [o["elapsed"] for r, o in enumerate(out) if r != straggler]
    ok = all(o["status"] == 0 and o["finite"] for o in out)
    if a.model > 0:
        paced = (a.steps - a.model - 1) * a.delay_ms / 1000.0 * 0.7
        ok = ok and all(o["lag"] <= a.model for o in out) and min(fast) >= paced
    else:
        ok = ok and max(fast) < 0.5 * total_delay
    print(f"MPSPARSE-BOUNDED model={a.model} world={world} fast ranks {min(fast):.3f}-{max(fast):.3f}s, straggler "
          f"{out[straggler]['elapsed']:.3f}s of {total_delay:.3f}s sleep, max lag {max(o['lag'] for o in out)} "
          f"loss {out[0]['first']:.4f} -> {out[0]['last']:.4f} -> {'OK' if ok else 'FAIL'}")
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
