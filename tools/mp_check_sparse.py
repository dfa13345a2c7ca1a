"""Multi-process check of the sharded sparse paths (direct P2P and slab exchange) against dense NumPy SGD.
torchrun launch, one rank per GPU; prints MPCHECK-SPARSE ... OK|FAIL on rank 0."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from psb200.data import random_csr
from psb200.engine.sparse_engine import DeviceCSR, SlabSparseEngine, SparseShardedEngine

p = argparse.ArgumentParser()
p.add_argument("--path", default="slab", choices=["direct", "slab"])
a = p.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
F, B, eta, steps = 6000, 256, 0.7, 4
eng = SlabSparseEngine(F, 5, rank, world, world, eta, B * 40, dev) if a.path == "slab" else \
    SparseShardedEngine(F, 5, rank, world, world, eta, 0, dev)
eng.connect()
losses = []
for step in range(steps):
    eng.step(DeviceCSR(random_csr(B, F, 40, seed=50 * step + rank), dev))
    torch.cuda.synchronize()
    losses.append(eng.mean_loss(DeviceCSR(random_csr(B, F, 40, seed=50 * step + rank), dev)))
dist.barrier()
shard = eng.shard_weights().cpu().numpy()
all_shards, all_losses = [None] * world, [None] * world
dist.all_gather_object(all_shards, shard)
dist.all_gather_object(all_losses, losses)
ok = eng.status() == 0
if rank == 0:
    W, b = np.zeros((F, 8)), np.zeros(8)
    ref = []
    for step in range(steps):
        dW, db, ls = np.zeros_like(W), np.zeros_like(b), []
        for r in range(world):
            bt = random_csr(B, F, 40, seed=50 * step + r)
            X = bt.to_dense(np.float64)
            z = X @ W[:, :6] + b[:6]
            z -= z.max(axis=1, keepdims=True)
            pr = np.exp(z)
            pr /= pr.sum(axis=1, keepdims=True)
            y = bt.labels.astype(np.int64)
            ls.append(float(-np.log(pr[np.arange(B), y]).mean()))
            pr[np.arange(B), y] -= 1
            dW[:, :6] += -eta / (B * world) * (X.T @ pr)
            db[:6] += -eta / (B * world) * pr.sum(axis=0)
        W, b = W + dW, b + db
        ref.append(ls)
    Fs = eng.Fs
    Wd = np.concatenate([s[:Fs] for s in all_shards])[:F]
    bd = all_shards[0][Fs]
    e_w = float(np.abs(Wd - W).max())
    e_b = float(np.abs(bd - b).max())
    e_l = float(np.abs(np.asarray(all_losses).T - np.asarray(ref)).max())
    ok = ok and e_w < 2e-4 * max(1.0, np.abs(W).max()) and e_b < 2e-4 and e_l < 3e-4
    print(f"MPCHECK-SPARSE path={a.path} world={world} w_err={e_w:.2e} bias_err={e_b:.2e} loss_err={e_l:.2e} -> "
          f"{'OK' if ok else 'FAIL'}")
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
