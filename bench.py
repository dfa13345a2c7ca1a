#!/usr/bin/env python
"""Headline benchmark: trained samples/sec of the parameter server, sequential consistency,
fine-food-reviews-shaped synthetic data, N B200s of one node (BASELINE.json config 2).

One "step" is one full parameter-server clock on every GPU: every worker pulls the current
weights, runs the reference's local solve (2 L-BFGS iterations with strong-Wolfe line search,
standardisation and centring — what Spark's LogisticRegression.setMaxIter(2).fit does in
LogisticRegressionTaskSpark.java:179-184) on its whole input buffer, pushes its delta, the
deltas are reduced and applied (w += delta/numWorkers, ServerProcessor.java:148-151) and the
new weights are published.  samples/sec = numWorkers x buffer rows x clocks / second — the same
accounting BASELINE.md uses for the reference (worker iterations/s x buffer size = ~130).

Two timed regions.  `value`: K clocks on resident buffers, CUDA events around the launches (device time).
`e2e`: the same K clocks through the public engine API with, every clock, 128 fresh rows per worker copied from
pinned host memory (packed: presence bitmap + per-row palette, bit-exact) and stored into the rings, and the
clock's per-worker losses copied back and read on the host.  The e2e loop is software-pipelined (copies of
batches i+1 and i+2 and the read of clock i-1's result overlap clock i) but nothing is skipped: bytes are
counted from the tensors copied, and its time is max(CUDA events, host wall clock).

Prints ONE JSON line on rank 0.  `--impl reference` reports that the Java/Kafka/Spark reference
cannot be installed here.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BASELINE_SAMPLES_PER_SEC = 130.0     # sequential, 4 workers (1.01 its/s x 128 rows)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--kernel", default="auto", choices=["auto", "simt", "tc"])
    p.add_argument("--workers_per_gpu", type=int, default=296, help="2 worker CTAs per SM")
    p.add_argument("--buffer", type=int, default=1024, help="rows per worker buffer (reference -max default)")
    p.add_argument("--features", type=int, default=1024)
    p.add_argument("--density", type=float, default=0.45, help="non-zero fraction of the hashed BoW rows")
    p.add_argument("--solver", default="lbfgs2", choices=["lbfgs2", "sgd"])
    p.add_argument("--consistency_model", type=int, default=0,
                   help="0 sequential (headline), k > 0 bounded delay, -1 eventual; the relaxed models run with per-worker "
                        "clocks (every worker gated on its own, its delta applied on arrival)")
    p.add_argument("--clocks_per_launch", type=int, default=1,
                   help="relaxed models: clocks every worker runs per launch at its own pace (a step is still ONE clock: "
                        "--steps must be a multiple)")
    p.add_argument("--ingest_rows", type=int, default=128, help="new rows per worker per step in the e2e run")
    p.add_argument("--ingest_format", choices=["palette", "bitmap", "dense"], default="palette",
                   help="e2e wire format of the fresh rows (all lossless w.r.t. the fp16 rings): presence bitmap + per-row "
                        "palette and 4-bit codes, presence bitmap + fp16 non-zeros, or dense fp16")
    p.add_argument("--no_e2e", action="store_true")
    p.add_argument("--stats", choices=["fused", "ingest"], default="fused",
                   help="standardisation statistics: recomputed by every step, fused into its first pass over the buffer "
                        "(what the reference's fit does), or kept current at ingestion time (windows that did not change "
                        "cost nothing)")
    p.add_argument("--no_overlap", action="store_true",
                   help="e2e run: store every batch between two steps instead of into spare ring slots while a step runs")
    p.add_argument("--events_per_min", type=float, default=0.0,
                   help="e2e run: per-worker arrival rate used for the dynamic buffer size bc x events/min (0: unthrottled -> max)")
    p.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                   help="p2p: fused peer-memory push/apply/pull (product); nccl: all_reduce baseline")
    return p.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for k, name in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def synth_buffers(torch, W, cap, F, density, seed, device):
    """Fine-food-shaped rows on the device: signed hashed counts at the given density, L2-normalised,
    labels 1..5 from a planted linear teacher plus noise (so the loss is meaningful)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    teacher = torch.randn(F, 5, generator=g, device=device)
    X = torch.empty((W, cap, F), dtype=torch.float16, device=device)
    y = torch.empty((W, cap), dtype=torch.int32, device=device)
    for w in range(W):
        mask = torch.rand((cap, F), generator=g, device=device) < density
        cnt = torch.randint(1, 4, (cap, F), generator=g, device=device).float()
        sgn = torch.randint(0, 2, (cap, F), generator=g, device=device).float() * 2 - 1
        v = mask.float() * cnt * sgn
        v = v / v.norm(dim=1, keepdim=True).clamp_min(1e-6)
        logits = v @ teacher * 4.0 + torch.randn((cap, 5), generator=g, device=device)
        X[w] = v.half()
        y[w] = (logits.argmax(dim=1) + 1).int()
    return X, y


def main():
    a = parse()
    if a.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:          # under torchrun: one line, from rank 0
            return 0
        print(json.dumps({"impl": "reference",
                          "unavailable": "reference is Java 8 + Gradle + Kafka + Spark (no setup.py/pyproject; pip "
                                         "install of /root/reference fails), and no JVM/Gradle/Kafka/Spark exists "
                                         "on this image or offline; published logs in BASELINE.md are the numbers "
                                         "of record"}))
        return 0
    import numpy as np
    import torch
    import torch.distributed as dist
    from psb200.config import PSConfig
    from psb200.engine.device_engine import DeviceEngine
    from psb200.ops import native

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus and world > 1:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from psb200.utils.numa import bind_to_device_node
    numa = bind_to_device_node(local)      # threads and pinned staging memory next to this rank's GPU (best effort,
    #                                        never raises; None on a single-node box or with PSB_NUMA_BIND=0)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG", "WARN")          # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    native.lib()
    impl = {"simt": native.IMPL_SIMT, "tc": native.IMPL_TC}.get(a.kernel)
    if impl is None:
        impl = native.IMPL_TC if os.environ.get("PSB_BENCH_TC", "1") == "1" else native.IMPL_SIMT
    W, cap, F = a.workers_per_gpu, a.buffer, a.features
    cfg = PSConfig(num_workers=W * world, num_features=F, consistency_model=a.consistency_model,
                   max_buffer_size=cap, min_buffer_size=min(128, cap), local_solver=a.solver)
    w0 = (np.random.default_rng(0).standard_normal((F + 1, 8)) * 0.01).astype(np.float32)   # random-init weights
    w0[:, 6:] = 0
    if a.transport == "nccl":
        from psb200.parallel.nccl_baseline import NcclBaselineEngine as Engine
    else:
        Engine = DeviceEngine
    kw = {} if a.no_overlap or a.no_e2e else {"ring_spare_rows": a.ingest_rows, "stats_at_ingest": a.stats == "ingest"}
    cpl = max(1, a.clocks_per_launch) if a.transport == "p2p" else 1
    if a.consistency_model != 0 and a.transport == "p2p":
        kw.update(async_granularity="worker", max_clocks_per_launch=cpl)
    if a.steps % cpl or a.warmup % cpl:
        raise SystemExit("--steps and --warmup must be multiples of --clocks_per_launch")
    eng = Engine(cfg, W, rank, world, impl, dev, w0=w0, stage_rows=W * a.ingest_rows, **kw)
    eng.connect()
    X, y = synth_buffers(torch, W, cap, F, a.density, 1234 + rank, dev)
    eng.fill(X, y)
    del X, y

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------------------------------------------------------- device-timed region
    def run_clocks(n):
        if cpl == 1:
            for _ in range(n):
                eng.step()
        else:
            for _ in range(n // cpl):
                eng.step(num_clocks=cpl)

    run_clocks(a.warmup)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = eng.kernel_launches
    e0.record()
    run_clocks(a.steps)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = eng.kernel_launches - launches0
    status = eng.status()
    loss = eng.read_loss().numpy().copy()
    info = eng.info.cpu().numpy()
    samples = W * world * cap * a.steps
    value = samples / (ms / 1000.0)

    # ---------------------------------------------------------------- end-to-end through the public API
    e2e = None
    if not a.no_e2e:
        n_new = W * a.ingest_rows
        g = torch.Generator().manual_seed(7 + rank)
        pool = 4                                              # rotating pinned batches of fresh rows
        # the same row model as the resident buffers (synth_buffers): signed hashed counts 1..3, L2-normalised
        host_rows = [(torch.rand((n_new, F), generator=g) < a.density).float()
                     * torch.randint(1, 4, (n_new, F), generator=g).float()
                     * (torch.randint(0, 2, (n_new, F), generator=g).float() * 2 - 1) for _ in range(pool)]
        host_rows = [(r / r.norm(dim=1, keepdim=True).clamp_min(1e-6)).half().pin_memory() for r in host_rows]
        host_lab = [torch.randint(1, 6, (n_new,), generator=g, dtype=torch.int32).pin_memory() for _ in range(pool)]
        if a.ingest_format != "dense":            # what the producer hands over: packed rows in pinned memory
            host_rows = [native.PackedRows(n_new, F, palette=a.ingest_format == "palette").pack_dense(r, l)
                         for r, l in zip(host_rows, host_lab)]
        workers = np.repeat(np.arange(W, dtype=np.int32), a.ingest_rows)
        # arrival timestamps drive the dynamic buffer size: target = clamp(bc * events_per_minute, min, max)
        gap_ms = 60000.0 / a.events_per_min if a.events_per_min > 0 else 0.0
        tick = [0]

        def stamps():
            base = tick[0] * a.ingest_rows * gap_ms
            tick[0] += 1
            return base + np.tile(np.arange(a.ingest_rows) * gap_ms, W)

        # Software pipeline through the public API: the host->device copies of batches i+2 and i+3
        # (ingest_begin, copy stream, three staging sets) are in flight while step i trains, and batch i+1 is stored
        # into the rings WHILE step i runs: the rings hold `ingest_rows` spare slots per worker, so the new rows land
        # outside the windows step i reads (ingest_commit(overlap=True)); step i+1 waits for that store.  Every step
        # still pays its own H2D (pinned host -> device) and its own D2H read of the losses.
        # The result read is pipelined the same way: step i's losses are copied device -> host behind step i and
        # consumed on the host after step i+1 has been queued (read_loss_begin / read_loss_end).
        pending = [None]
        seen = [0.0]
        overlap = not a.no_overlap

        def one(i):
            eng.step()                                                                     # enqueue clock i
            ticket = eng.read_loss_begin()                                                 # D2H of clock i's result
            eng.ingest_commit(overlap=overlap)                                             # batch i+1 -> rings
            eng.ingest_begin(host_rows[(i + 3) % pool], host_lab[(i + 3) % pool], workers, stamps())   # H2D of i+3
            if pending[0] is not None:
                seen[0] = float(eng.read_loss_end(pending[0])[0])                          # clock i-1's result, on the host
            pending[0] = ticket

        def drain():
            if pending[0] is not None:
                seen[0] = float(eng.read_loss_end(pending[0])[0])
                pending[0] = None

        eng.ingest_begin(host_rows[0], host_lab[0], workers, stamps())
        eng.ingest_commit()                                              # batch 0 before the first step
        eng.ingest_begin(host_rows[1], host_lab[1], workers, stamps())
        eng.ingest_begin(host_rows[2], host_lab[2], workers, stamps())
        for i in range(a.warmup):
            one(i)
        drain()
        barrier()
        h0 = eng.h2d_bytes
        t0 = torch.cuda.Event(enable_timing=True)
        t1 = torch.cuda.Event(enable_timing=True)
        t0.record()
        wall0 = time.perf_counter()
        for i in range(a.steps):
            one(i)
        drain()                                                          # the last clock's result is read inside the timed region
        t1.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - wall0) * 1000.0
        barrier()
        ems = max_over_ranks(max(t0.elapsed_time(t1), wall))
        eng.ingest_commit(), eng.ingest_commit()                       # drain the two batches still in the pipeline
        rows_eff = float(eng.win_host[:, 1].float().mean())            # dynamic buffer size actually trained on
        e2e = {"value": W * world * rows_eff * a.steps / (ems / 1000.0), "unit": "samples/sec",
               "buffer_rows_effective": rows_eff, "events_per_min_per_worker": a.events_per_min,
               "h2d_bytes_per_step": (eng.h2d_bytes - h0) // a.steps, "d2h_bytes_per_step": W * 4,
               "ms_per_step": ems / a.steps, "ingested_samples_per_sec": n_new * world * a.steps / (ems / 1000.0),
               "new_rows_per_worker_per_step": a.ingest_rows, "ingest_format": a.ingest_format,
               "store_overlaps_step": bool(overlap and eng.ring_spare >= a.ingest_rows),
               "e2e_check": {"inputs_from_pinned_host_every_step": True, "h2d_inside_timed_region": True,
                             "result_read_on_host_every_step": True, "d2h_inside_timed_region": True,
                             "timed_by": "max(CUDA events, host wall clock), max over ranks"}}

    clocks = sampler.stop() if rank == 0 else None       # sampled across both timed regions
    if rank == 0:
        out = {
            "metric": "samples/sec (trained samples: numWorkers x buffer rows x clocks/s, sequential consistency)",
            "value": value, "unit": "samples/sec", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": value / BASELINE_SAMPLES_PER_SEC, "dtype": "fp16 inputs, fp32 weights fed to tcgen05 as fp16 hi+lo pairs, fp32 accumulate",
            "data": "synthetic (fine-food-reviews-shaped hashed BoW rows generated on device; random-init weights)",
            "impl": "ours", "transport": a.transport, "kernel": "tcgen05" if impl == native.IMPL_TC else "simt",
            "gpu_launches": launches, "clocks": clocks, "e2e": e2e,
            "config": {"model": f"multinomial logistic regression {F}->6 class rows ({(F + 1) * 6} params)",
                       "global_batch": W * world * cap, "seq_len": None,
                       "parallelism": f"ps-dp{world} ({W} workers/GPU, {world} server shards)",
                       "num_workers": W * world, "buffer_rows": cap, "density": a.density,
                       "local_solver": a.solver, "consistency_model": a.consistency_model, "statistics": a.stats,
                       "clocks_per_launch": cpl, "numa": numa,
                       "l2": f"inputs larger than L2 ({W * cap * F * 2 / 2**20:.0f} MiB of rows per GPU streamed per pass)",
                       "evals_per_step": float(info[:, 0].mean()), "status": status,
                       "cycles64_stats_solve_push_total": [float(info[:, k].mean()) for k in (4, 5, 6, 7)],
                       "mean_loss": float(np.nanmean(loss))},
        }
        print(json.dumps(out), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
