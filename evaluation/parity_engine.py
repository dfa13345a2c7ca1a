"""F1 parity of the PRODUCT path per consistency model: `psb200.apps.node` (DeviceEngine: fused kernels, peer-memory
transport, one process per GPU) against the CPU oracle on the same synthetic stream.

  python evaluation/parity_engine.py --oracle            CPU: lock-step oracle (fp64 L-BFGS restatement of the
                                                         reference's local solver), cached in logs/parity_engine_oracle.npz
  python evaluation/parity_engine.py --gpus 8            GPU box: runs the node app under torchrun for -c 0 / 10 / -1,
                                                         keeps the raw server logs in logs/parity_engine_<model>_server.csv
                                                         and prints the comparison table

Stream: row i goes to worker i % numWorkers (CsvProducer.java:61); with -p 0 every worker ingests 128 rows per clock
and trains on its last <= 1024 rows, so the sequential model is deterministic and is compared clock by clock; the relaxed
models are compared through the mean F1 of their last rows (BASELINE.json: within 0.5 pt per consistency model)."""
import argparse
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
import numpy as np

ORACLE = os.path.join(HERE, "logs", "parity_engine_oracle.npz")
MODELS = ((0, "sequential"), (10, "bounded delay 10"), (-1, "eventual"))


def oracle(workers: int, clocks: int, rows_per_clock: int, cap: int, test_rows: int):
    from psb200.config import PSConfig
    from psb200.data import CSRBatch, ReviewGenerator
    from psb200.models import LogRegLayout
    from psb200.oracle import LogisticRegressionTask
    cfg = PSConfig(num_workers=workers)
    gen = ReviewGenerator(cfg.num_features, cfg.num_classes, seed=cfg.seed)
    n = workers * rows_per_clock * clocks
    train, test = gen.train_split(n, seed=cfg.seed + 1), gen.test_split(test_rows, seed=cfg.seed + 2)
    half = lambda b: CSRBatch.from_dense(b.to_dense(np.float16).astype(np.float32), b.labels)
    train, test = half(train), half(test)
    mine = [np.arange(w, n, workers) for w in range(workers)]
    lay = LogRegLayout(cfg.num_features, cfg.num_classes)
    W = lay.zeros()
    scorer = LogisticRegressionTask(cfg.num_features, cfg.num_classes, test, "lbfgs2")
    f1s, losses = [], []
    for v in range(clocks):
        tot, ls = np.zeros_like(W, dtype=np.float64), []
        for w in range(workers):
            seen = rows_per_clock * (v + 1)
            rows = mine[w][max(0, seen - cap):seen]
            t = LogisticRegressionTask(cfg.num_features, cfg.num_classes, None, "lbfgs2")
            t.set_weights(W)
            tot += t.calculate_gradients(train.slice(rows), eval_test=False)
            ls.append(t.loss)
        W = (W + tot / workers).astype(np.float32)
        scorer.set_weights(W)
        m = scorer.calculate_test_metrics()
        f1s.append(m.f1), losses.append(float(np.mean(ls)))
        print(f"oracle clock {v}: F1 {m.f1:.4f} loss {losses[-1]:.4f}", flush=True)
    return np.asarray(f1s), np.asarray(losses)


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--oracle", action="store_true")
    p.add_argument("--gpus", type=int, default=0)
    p.add_argument("--workers_per_gpu", type=int, default=2)
    p.add_argument("--total_workers", type=int, default=16)
    p.add_argument("--clocks", type=int, default=60)
    p.add_argument("--test_rows", type=int, default=4877)
    a = p.parse_args()
    if a.oracle:
        f1, loss = oracle(a.total_workers, a.clocks, 128, 1024, a.test_rows)
        np.savez(ORACLE, f1=f1, loss=loss, workers=a.total_workers, clocks=a.clocks)
        print(f"oracle: F1 mean of the last 20 clocks {f1[-20:].mean():.4f}; saved {ORACLE}")
        return 0
    from psb200.utils.logs import parse_log
    ref = np.load(ORACLE)
    assert a.gpus * a.workers_per_gpu == int(ref["workers"]), "run the oracle with the same number of workers"
    n = int(ref["workers"]) * 128 * int(ref["clocks"])
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    print("| consistency | clocks | oracle F1 (mean last 20) | engine F1 (mean last 20) | diff, pt | max diff over all clocks, pt |")
    print("|---|---|---|---|---|---|")
    for model, name in MODELS:
        tmp = tempfile.mkdtemp(prefix="psb_parity_")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(29600 + (model + 2) % 50), "-m", "psb200.apps.node",
               "-training", f"synthetic:{n}", "-test", f"synthetic:{a.test_rows}", "-p", "0", "--workers_per_gpu",
               str(a.workers_per_gpu), "-c", str(model), "--max_clock", str(int(ref["clocks"])), "-l",
               "--clocks_per_launch", "1"]
        r = subprocess.run(cmd, env=env, cwd=tmp, capture_output=True, text=True, timeout=1800)
        if r.returncode != 0:
            print(f"| {name} | node app failed | | | | |")
            print(r.stderr[-2000:], file=sys.stderr)
            continue
        tag = {0: "sequential", 10: "bounded10", -1: "eventual"}[model]
        for role in ("server", "worker"):
            shutil.copy(os.path.join(tmp, f"logs-{role}.csv"), os.path.join(HERE, "logs", f"parity_engine_{tag}_{role}.csv"))
            if os.path.isdir(os.path.join(ROOT, "gpurun_out")):       # on a gpurun box only gpurun_out/ travels back
                shutil.copy(os.path.join(tmp, f"logs-{role}.csv"),
                            os.path.join(ROOT, "gpurun_out", f"parity_engine_{tag}_{role}.csv"))
        s = parse_log(os.path.join(tmp, "logs-server.csv"), "server")
        k = min(len(s), len(ref["f1"]))
        d = np.abs(s[:k, 4] - ref["f1"][:k])
        per_clock = f"{100 * d.max():.3f}" if model == 0 else "(asynchronous)"
        print(f"| {name} | {len(s)} | {ref['f1'][-20:].mean():.4f} | {s[-20:, 4].mean():.4f} | "
              f"{100 * abs(s[-20:, 4].mean() - ref['f1'][-20:].mean()):.3f} | {per_clock} |")
        shutil.rmtree(tmp, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
