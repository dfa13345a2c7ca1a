"""Multi-process check of the peer-memory protocol: one rank per GPU under torchrun.
Every rank steps the fused kernel for a few clocks, then rank 0 compares the replicated weights and the
per-worker losses with the lock-step CPU oracle.  Usage (2 GPUs):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
      tools/mp_check.py --model 0 --impl 1
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from psb200.config import PSConfig
from psb200.data import CSRBatch, ReviewGenerator
from psb200.engine.device_engine import DeviceEngine
from psb200.models import LogRegLayout
from psb200.oracle import LogisticRegressionTask

p = argparse.ArgumentParser()
p.add_argument("--model", type=int, default=0)
p.add_argument("--impl", type=int, default=1)
p.add_argument("--wpr", type=int, default=2)
p.add_argument("--clocks", type=int, default=4)
a = p.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
F, cap, n = 256, 128, 96
cfg = PSConfig(num_workers=world * a.wpr, num_features=F, consistency_model=a.model, max_buffer_size=cap,
               min_buffer_size=16, local_solver="lbfgs2")
gen = ReviewGenerator(F, 5, vocab=3000, mean_len=120.0)
full = gen.train_split(n * world * a.wpr)
full = CSRBatch.from_dense(full.to_dense(np.float16).astype(np.float32), full.labels)
per_worker = [full.slice(np.arange(i * n, (i + 1) * n)) for i in range(world * a.wpr)]
eng = DeviceEngine(cfg, a.wpr, rank, world, a.impl, dev)
eng.connect()
mine = per_worker[rank * a.wpr:(rank + 1) * a.wpr]
eng.fill(torch.stack([torch.from_numpy(b.to_dense(np.float16)) for b in mine]).to(dev),
         torch.stack([torch.from_numpy(b.labels) for b in mine]).to(dev))
dist.barrier()
losses = []
for _ in range(a.clocks):
    eng.step()
    losses.append(eng.read_loss().numpy().copy())
torch.cuda.synchronize()
dist.barrier()
status = eng.status()
all_losses = [None] * world
dist.all_gather_object(all_losses, np.asarray(losses))
ok = status == 0
if a.model == 0:
    W_dev = eng.weights().cpu().numpy()
    all_w = [None] * world
    dist.all_gather_object(all_w, W_dev)
    if rank == 0:
        lay = LogRegLayout(F, 5)
        Wref = lay.zeros()
        lref = []
        for _ in range(a.clocks):
            tot = np.zeros_like(Wref, dtype=np.float64)
            ls = []
            for b in per_worker:
                t = LogisticRegressionTask(F, 5, None, "lbfgs2")
                t.set_weights(Wref)
                tot += t.calculate_gradients(b, eval_test=False)
                ls.append(t.loss)
            Wref = (Wref + tot / len(per_worker)).astype(np.float32)
            lref.append(ls)
        got = np.concatenate(all_losses, axis=1)
        e_loss = float(np.abs(got - np.asarray(lref)).max())
        e_w = max(float(np.abs(w - Wref).max()) for w in all_w)
        same = all(np.array_equal(all_w[0], w) for w in all_w)
        ok = ok and e_loss < 5e-3 and e_w < 5e-3 * max(1e-2, float(np.abs(Wref).max())) and same
        print(f"MPCHECK model={a.model} impl={a.impl} world={world} loss_err={e_loss:.2e} w_err={e_w:.2e} "
              f"replicas_identical={same} status={status} -> {'OK' if ok else 'FAIL'}")
else:
    # async modes: ranks free-run; check liveness, finite losses, decreasing loss and the clock table
    got = np.concatenate(all_losses, axis=1)
    fin = bool(np.isfinite(got).all())
    dec = bool(got[-1].mean() < got[0].mean())
    ok = ok and fin and dec
    if rank == 0:
        print(f"MPCHECK model={a.model} impl={a.impl} world={world} finite={fin} loss {got[0].mean():.4f} -> "
              f"{got[-1].mean():.4f} status={status} -> {'OK' if ok else 'FAIL'}")
eng.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
