"""Multi-process check of the per-worker consistency models on real GPUs (one rank per GPU under torchrun).

Every rank launches ITS workers for --clocks clocks in ONE launch (per-worker clocks, PSB_COMM_WORKER); global worker
--straggler sleeps before each of its clocks.  The gates are evaluated inside the kernels on the clock table that every
worker replicates into every rank's block, so what is checked is the reference's contract
(MessageTracker.java:69-87): bounded delay k -> nobody starts a clock more than k ahead of the slowest worker, and the
fast ones do reach exactly k; eventual -> nobody waits.  Usage (2 GPUs):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      tools/mp_worker_clocks.py --model 2
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from psb200.config import PSConfig
from psb200.data import ReviewGenerator
from psb200.engine.device_engine import DeviceEngine

p = argparse.ArgumentParser()
p.add_argument("--model", type=int, default=2)
p.add_argument("--impl", type=int, default=1)
p.add_argument("--wpr", type=int, default=2)
p.add_argument("--clocks", type=int, default=12)
p.add_argument("--straggler", type=int, default=-2, help="global worker id (-2: first worker of the last rank)")
p.add_argument("--delay_cycles", type=int, default=3_000_000)
a = p.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
F, cap, n = 256, 128, 96
cfg = PSConfig(num_workers=world * a.wpr, num_features=F, consistency_model=a.model, max_buffer_size=cap,
               min_buffer_size=16)
eng = DeviceEngine(cfg, a.wpr, rank, world, a.impl, dev, async_granularity="worker", max_clocks_per_launch=a.clocks,
                   spin_limit_cycles=8_000_000_000)
eng.connect()
b = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0).train_split(n * a.wpr, seed=20 + rank)
eng.fill(torch.from_numpy(b.to_dense(np.float16)).view(a.wpr, n, F).to(dev),
         torch.from_numpy(b.labels.astype(np.int32)).view(a.wpr, n).to(dev))
straggler = (world - 1) * a.wpr if a.straggler == -2 else a.straggler
eng.inject_delay(straggler, a.delay_cycles)
dist.barrier()
eng.step(num_clocks=a.clocks)
h = eng.read_hist()
torch.cuda.synchronize()
dist.barrier()
table = eng.worker_clocks()
W = eng.weights().cpu().numpy()
gathered = [None] * world
dist.all_gather_object(gathered, {"lag": h["lag"], "lead": h["lead"], "loss": h["loss"], "status": eng.status(),
                                  "table": table, "w_finite": bool(np.isfinite(W).all())})
ok = True
if rank == 0:
    lag = np.concatenate([g["lag"] for g in gathered], axis=1)          # [clocks, all workers]
    loss = np.concatenate([g["loss"] for g in gathered], axis=1)
    fast = np.delete(lag, straggler, axis=1)
    ok = all(g["status"] == 0 and g["w_finite"] and (g["table"] == a.clocks).all() for g in gathered)
    ok = ok and bool(np.isfinite(loss).all())
    if a.model > 0:
        ok = ok and int(lag.max()) == a.model and int(fast.max()) == a.model and int(lag[:, straggler].max()) <= 1
    else:
        ok = ok and int(fast.max()) >= a.clocks - 3
    print(f"MPWORKERCLOCKS model={a.model} impl={a.impl} world={world} max_lag={int(lag.max())} "
          f"straggler_lag={int(lag[:, straggler].max())} loss {loss[0].mean():.4f} -> {loss[-1].mean():.4f} "
          f"-> {'OK' if ok else 'FAIL'}")
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
