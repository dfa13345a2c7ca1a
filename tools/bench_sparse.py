"""BASELINE config 5: 2^24-feature hashed sparse LR, server shards across the GPUs — push/pull bandwidth
of the fused P2P path versus an NCCL data-parallel baseline (replicated model, dense all_reduce of the
gradient).  Launch like bench.py (torchrun for N > 1).  One JSON line per configuration on rank 0."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from psb200.engine.sparse_engine import SlabSparseEngine, SparseShardedEngine
from psb200.ops import native


class Batch:
    def __init__(self, B, F, nnz, seed, dev):
        g = torch.Generator(device=dev).manual_seed(seed)
        idx = torch.randint(0, F, (B, nnz), generator=g, device=dev, dtype=torch.int64).sort(dim=1).values
        val = torch.randn((B, nnz), generator=g, device=dev)
        val = val / val.norm(dim=1, keepdim=True)
        self.indptr = (torch.arange(B + 1, device=dev, dtype=torch.int64) * nnz)
        self.indices = idx.reshape(-1).int().contiguous()
        self.values = val.reshape(-1).contiguous()
        self.labels = torch.randint(1, 6, (B,), generator=g, device=dev, dtype=torch.int32)
        self.n, self.nnz = B, B * nnz


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--features", type=int, default=1 << 24)
    p.add_argument("--batch", type=int, default=16384)
    p.add_argument("--nnz", type=int, default=128)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--consistency_model", type=int, default=-1)
    p.add_argument("--path", default="direct", choices=["direct", "slab"],
                   help="direct: one P2P request per non-zero; slab: request lists + contiguous reply/update slabs")
    p.add_argument("--l2_fetch", type=int, default=0, help="cudaLimitMaxL2FetchGranularity hint (32/64/128; 0 = leave)")
    p.add_argument("--baseline", action="store_true", help="also time the NCCL dense all_reduce baseline")
    a = p.parse_args()
    world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    F = a.features
    l2 = native.lib().psb_set_l2_fetch_granularity(a.l2_fetch) if a.l2_fetch else None
    batches = [Batch(a.batch, F, a.nnz, 100 * rank + i, dev) for i in range(4)]

    def timed(fn):
        for i in range(a.warmup):
            fn(i)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.steps):
            fn(i)
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / a.steps

    if a.path == "slab":
        eng = SlabSparseEngine(F, 5, rank, world, world, 0.5, a.batch * a.nnz, dev)
    else:
        eng = SparseShardedEngine(F, 5, rank, world, world, 0.5, a.consistency_model, dev)
    eng.connect()
    ms = timed(lambda i: eng.step(batches[i % 4]))
    loss = eng.mean_loss(batches[(a.steps - 1) % 4])
    status = eng.status()
    bytes_pp = 2 * batches[0].nnz * 32                         # pull + push bytes per GPU per step
    out = {"config": "sharded sparse LR", "impl": "fused-p2p-" + a.path, "n_gpus": world, "features": F, "batch_per_gpu": a.batch,
           "nnz_per_row": a.nnz, "consistency_model": a.consistency_model, "ms_per_step": ms,
           "samples_per_sec": a.batch * world / (ms / 1e3),
           "pull_push_GBps_per_gpu": bytes_pp / (ms / 1e3) / 1e9,
           "nvlink_GBps_per_gpu_per_direction": bytes_pp / 2 * (world - 1) / world / (ms / 1e3) / 1e9,
           "model_MiB_per_shard": eng.shard_bytes / 2 ** 20, "loss": loss, "status": status,
           "l2_fetch_granularity": l2, "kernels_per_step": 4 if a.path == "slab" else (1 if a.consistency_model != 0 else 2)}
    if rank == 0:
        print(json.dumps(out))
    eng.close()

    if a.baseline:
        # NCCL data-parallel baseline: replicated model, local sparse gradient into a dense buffer,
        # all_reduce of the dense gradient (512 MiB at 2^24 features), dense apply.
        rep = SparseShardedEngine(F, 5, 0, 1, world, 0.5, 0, dev)          # full replica on every GPU
        grad = SparseShardedEngine(F, 5, 0, 1, world, 0.5, 0, dev)         # dense gradient buffer, same layout
        from psb200.engine.device_engine import _view
        Wv = _view(rep.shard, (rep.Fs + 1) * 8, dev)
        Gv = _view(grad.shard, (grad.Fs + 1) * 8, dev)

        def base_step(i):
            b = batches[i % 4]
            rep.loss.zero_()
            rep._launch(1, 0, 0, b)                                         # forward from the replica
            grad.residuals = rep.residuals
            grad._launch(2, 0, 0, b)                                        # scatter into the dense gradient
            if world > 1:
                dist.all_reduce(Gv)
            Wv.add_(Gv)
            Gv.zero_()

        ms_b = timed(base_step)
        if rank == 0:
            print(json.dumps({"config": "sharded sparse LR", "impl": "nccl-dense-allreduce-baseline", "n_gpus": world,
                              "features": F, "batch_per_gpu": a.batch, "nnz_per_row": a.nnz, "ms_per_step": ms_b,
                              "samples_per_sec": a.batch * world / (ms_b / 1e3),
                              "allreduce_MiB": (rep.Fs + 1) * 32 / 2 ** 20, "speedup_of_fused": ms_b / ms}))
        rep.close(), grad.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
