"""Smallest workload touching every kernel once, for compute-sanitizer (one tool per gpurun call)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from psb200.config import PSConfig
from psb200.data import ReviewGenerator, random_csr
from psb200.engine.device_engine import VirtualCluster
from psb200.engine.sparse_engine import DeviceCSR, VirtualSparseCluster

F, cap, W = 256, 128, 2
gen = ReviewGenerator(F, 5, vocab=2000, mean_len=100.0)
b = gen.train_split(2 * W * cap)
te = gen.test_split(300)
for impl in (1, 0):
    for model in (0, -1):
        cfg = PSConfig(num_workers=2 * W, num_features=F, consistency_model=model, max_buffer_size=cap, min_buffer_size=16)
        vc = VirtualCluster(cfg, 2, W, impl)
        for r, e in enumerate(vc.engines):
            rows = torch.from_numpy(b.to_dense(np.float16)).view(2, W, cap, F)[r].cuda()
            labs = torch.from_numpy(b.labels).view(2, W, cap)[r].cuda()
            e.fill(rows, labs)
            e.load_test(te)
        for _ in range(2):
            vc.step()
        pin = torch.from_numpy(b.to_dense(np.float16)[:64]).pin_memory()
        vc.engines[0].ingest(pin, torch.from_numpy(b.labels[:64].astype(np.int32)).pin_memory(),
                             np.arange(64, dtype=np.int32) % W, np.zeros(64))
        vc.step()
        if model == 0:
            print("impl", impl, "f1/acc", vc.engines[0].evaluate(), "status", [e.status() for e in vc.engines])
        vc.close()
sc = VirtualSparseCluster(4096, 2, 0)
batches = [DeviceCSR(random_csr(128, 4096, 20, seed=s), sc.engines[0].device) for s in range(2)]
sc.step(batches)
print("sparse", sc.engines[0].mean_loss(batches[0]), sc.engines[0].evaluate(batches[0]))
sc.close()
torch.cuda.synchronize()
print("SANITIZE TARGET DONE")
