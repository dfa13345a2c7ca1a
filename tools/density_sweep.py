"""Dense fp16 rings + tcgen05 kernel vs sparse-row (ELL) rings + psb_worker_step_csr, per row density.
One clock = every worker runs the reference's 2-iteration L-BFGS local solve on its whole 1024-row buffer (1 GPU,
sequential protocol).  Usage: python tools/density_sweep.py [--workers 296]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from psb200.config import PSConfig
from psb200.engine.device_engine import DeviceEngine
from psb200.engine.sparse_row_engine import SparseRowEngine
from psb200.ops import native

p = argparse.ArgumentParser()
p.add_argument("--workers", type=int, default=296)
p.add_argument("--steps", type=int, default=20)
a = p.parse_args()
dev = torch.device("cuda", 0)
cap = 1024


def rows(W, F, density, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    X = torch.empty((W, cap, F), dtype=torch.float16, device=dev)
    y = torch.empty((W, cap), dtype=torch.int32, device=dev)
    teacher = torch.randn(F, 5, generator=g, device=dev)
    for w in range(W):
        v = (torch.rand((cap, F), generator=g, device=dev) < density).float() \
            * torch.randint(1, 4, (cap, F), generator=g, device=dev).float() \
            * (torch.randint(0, 2, (cap, F), generator=g, device=dev).float() * 2 - 1)
        v = v / v.norm(dim=1, keepdim=True).clamp_min(1e-6)
        X[w] = v.half()
        y[w] = ((v @ teacher) * 4 + torch.randn((cap, 5), generator=g, device=dev)).argmax(dim=1).int() + 1
    return X, y


def timed(eng, steps):
    for _ in range(3):
        eng.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


out = []
for F, density, W in ((1024, 0.45, a.workers), (1024, 0.25, a.workers), (1024, 0.05, a.workers), (4096, 0.05, 148),
                      (16384, 0.01, 148)):
    X, y = rows(W, F, density, 1)
    cfg = PSConfig(num_workers=W, num_features=F, max_buffer_size=cap)
    rec = {"features": F, "density": density, "workers": W, "buffer_rows": cap}
    if F <= 1024:
        eng = DeviceEngine(cfg, W, impl=native.IMPL_TC)
        eng.fill(X, y)
        ms = timed(eng, a.steps)
        rec["dense_tcgen05_ms"], rec["dense_loss"] = ms, float(eng.read_loss().mean())
        rec["dense_bytes_per_row"] = 2 * F
        eng.close()
    # the same rows as ELL rings
    nz = (X != 0)
    K = int(nz.sum(dim=2).max())
    K = (K + 31) // 32 * 32
    eng = SparseRowEngine(cfg, W, K)
    order = torch.argsort(nz.to(torch.int8), dim=2, descending=True, stable=True)[:, :, :K]     # non-zero columns first
    eng.idx.copy_(order.to(torch.int32))
    eng.val.copy_(torch.gather(X, 2, order))
    eng.nnz.copy_(nz.sum(dim=2).to(torch.int32))
    eng.labels.copy_(y)
    eng.win_host[:, 0], eng.win_host[:, 1] = 0, cap
    eng.win.copy_(eng.win_host)
    ms = timed(eng, a.steps)
    rec["sparse_rows_ms"], rec["sparse_loss"], rec["ell_width"] = ms, float(eng.read_loss().mean()), K
    rec["sparse_bytes_per_row"] = 6 * float(nz.sum(dim=2).float().mean())
    eng.close()
    rec["samples_per_s_best"] = W * cap / (min(v for k, v in rec.items() if k.endswith("_ms")) / 1000.0)
    out.append(rec)
    print(json.dumps(rec), flush=True)
    del X, y, nz, order
