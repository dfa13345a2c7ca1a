"""F1 parity per consistency model: the sm_100a kernels (CudaBackend, tcgen05 flavour) against the CPU oracle
(NumpyBackend) on the SAME synthetic stream, same virtual clock, same straggler pattern.  BASELINE.json's
target is "test F1 within 0.5 pt per consistency model"; the real review data are not available, so parity
is judged new build vs oracle.  Prints a markdown table; needs a GPU.

`--oracle_only` runs the CPU oracle arms and stores their server logs in evaluation/logs/parity_oracle.npz (no
GPU needed); a later run on a GPU box reuses that file instead of repeating the CPU minutes there."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from psb200.config import PSConfig
from psb200.data import CSRBatch, ReviewGenerator
from psb200.engine.cuda_backend import CudaBackend
from psb200.engine.sim import NumpyBackend, PSSimulation
from psb200.utils.logs import parse_log


def iter_time(worker, clock):
    jitter = ((worker * 7919 + clock * 104729) % 1000) / 1000.0 - 0.5
    return 2700.0 * (1.0 + 0.03 * worker) * (1.0 + 0.10 * jitter)


ORACLE_CACHE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "logs", "parity_oracle.npz")


def main():
    oracle_only = "--oracle_only" in sys.argv
    cache = dict(np.load(ORACLE_CACHE)) if os.path.exists(ORACLE_CACHE) and not oracle_only else {}
    gen = ReviewGenerator(mean_len=420.0)
    half = lambda b: CSRBatch.from_dense(b.to_dense(np.float16).astype(np.float32), b.labels)
    train, test = half(gen.train_split(6000)), half(gen.test_split())
    print("| consistency | clocks | oracle F1 (mean last 50) | sm_100a F1 (mean last 50) | |diff| pt | max |diff| over all clocks, pt |")
    print("|---|---|---|---|---|---|")
    for model, name in ((0, "sequential"), (10, "bounded delay 10"), (-1, "eventual")):
        logs = []
        for impl in ("numpy",) if oracle_only else ("numpy", 1, 0):
            if impl == "numpy" and name in cache:
                logs.append(cache[name])
                continue
            cfg = PSConfig(num_workers=4, consistency_model=model, producer_time_per_event=100)
            backend = NumpyBackend(cfg, test) if impl == "numpy" else CudaBackend(cfg, test, impl)
            sim = PSSimulation(cfg, train, test, backend=backend, iter_time_ms=iter_time, virtual=True,
                               worker_metrics=False)
            sim.run(max_clock=150)
            logs.append(parse_log(sim.server_log.rows, "server"))
        if oracle_only:
            cache[name] = logs[0]
            np.savez(ORACLE_CACHE, **cache)
            print(f"| {name} (oracle only) | {len(logs[0])} | {logs[0][-50:, 4].mean():.4f} | | | |")
            continue
        o, t, s = logs
        n = min(len(o), len(t), len(s))
        for label, g in (("tcgen05", t), ("simt", s)):
            d = np.abs(o[:n, 4] - g[:n, 4])
            print(f"| {name} ({label}) | {n} | {o[-50:, 4].mean():.4f} | {g[-50:, 4].mean():.4f} | "
                  f"{100 * abs(o[-50:, 4].mean() - g[-50:, 4].mean()):.3f} | {100 * d.max():.3f} |")


if __name__ == "__main__":
    main()
