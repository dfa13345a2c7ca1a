"""Fused multi-GPU entry point: ``torchrun --nproc-per-node N -m psb200.apps.node [flags]``.

One process per GPU; every rank hosts a server shard, ``--workers_per_gpu`` workers and its slice of
the producer.  It accepts the union of the ServerAppRunner and WorkerAppRunner options
(``ServerAppRunner.java:19-26``, ``WorkerAppRunner.java:17-24``) with their defaults, writes the same
``logs-server.csv`` (rank 0) and ``logs-worker.csv`` (``logs-worker-<rank>.csv`` beyond rank 0) rows,
and runs the streaming loop on the device engine: producer slice -> pinned staging -> ring scatter ->
one fused step kernel per clock -> evaluation on rank 0.

Extras over the reference: ``--checkpoint_dir/--checkpoint_every/--resume``, ``--inject_delay rank:ms``
(straggler experiments, README 5.3), ``--watchdog_s``.
"""
from __future__ import annotations

import os
import sys
import time
from typing import List, Optional

import numpy as np

from ..config import SERVER_LOG_HEADER, WORKER_LOG_HEADER
from ..utils.logs import LogSink, server_row, worker_row
from . import cli
from .base_app import load_csv_or_synthetic


def build_parser():
    p = cli.server_parser()
    p.prog = "psb200.apps.node"
    p.add_argument("--workers_per_gpu", type=int, default=1)
    p.add_argument("--impl", choices=["tc", "simt"], default="tc")
    p.add_argument("--rows_per_clock", type=int, default=0,
                   help="rows each worker ingests per clock when -p 0 (default: min buffer size)")
    p.add_argument("--checkpoint_dir", default="")
    p.add_argument("--checkpoint_every", type=int, default=0)
    p.add_argument("--resume", action="store_true")
    p.add_argument("--watchdog_s", type=float, default=120.0)
    p.add_argument("--worker_metrics", action="store_true", help="evaluate every worker's local model (slow)")
    return p


def main(argv: Optional[List[str]] = None) -> int:
    ns = cli.parse(build_parser(), argv)
    from ..utils.logging_config import configure
    configure()
    import torch
    import torch.distributed as dist
    from ..engine.device_engine import DeviceEngine
    from ..ops import native
    from ..utils import checkpoint
    from ..utils.watchdog import Watchdog

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    W = ns.workers_per_gpu
    ns.num_workers = W * world
    cfg = cli.config_from_args(ns, "server")
    if cfg.verbose and rank == 0:
        print("\nUsed parameter:")
        for k in ("training_data_file_path", "test_data_file_path", "consistency_model", "producer_time_per_event",
                  "min_buffer_size", "max_buffer_size", "buffer_size_coefficient", "num_workers"):
            print(f"    {k}: {getattr(cfg, k)}")
        print(f"    ranks: {world} x {W} workers\n")
    streaming_file = not cfg.training_data_file_path.startswith("synthetic:")
    train = None if streaming_file else load_csv_or_synthetic(cfg.training_data_file_path, cfg, "train")
    test = load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test")
    impl = native.IMPL_TC if ns.impl == "tc" else native.IMPL_SIMT
    if impl == native.IMPL_TC and (cfg.num_features % 128 or cfg.num_features > 1024 or cfg.max_buffer_size % 128):
        impl = native.IMPL_SIMT
    w0, resume_meta = None, None
    from ..models.logreg import LogRegLayout
    layout = LogRegLayout(cfg.num_features, cfg.num_classes)
    if ns.resume and ns.checkpoint_dir:
        w0, resume_meta = checkpoint.load(ns.checkpoint_dir, layout)
    elif cfg.init == "random":
        w0 = layout.random(cfg.seed)
    rows_per_clock = ns.rows_per_clock or cfg.min_buffer_size
    eng = DeviceEngine(cfg, W, rank, world, impl, dev, w0=w0, stage_rows=max(W * rows_per_clock, 1024))
    eng.connect()
    if resume_meta:
        eng.clock = int(resume_meta["clock"])
        if world > 1 or True:
            # version / clock words restart at the resumed clock on every rank
            eng.reset_clock_words(eng.clock)
    eng.load_test(test)
    server_log = worker_log = None
    if rank == 0:
        server_log = LogSink.to_file("./logs-server.csv", SERVER_LOG_HEADER) if cfg.logging else LogSink()
    wl_name = "./logs-worker.csv" if rank == 0 else f"./logs-worker-{rank}.csv"
    worker_log = LogSink.to_file(wl_name, WORKER_LOG_HEADER) if cfg.logging else (LogSink() if rank == 0 else None)

    # this rank's slice of the stream: global row i goes to worker i % numWorkers (CsvProducer.java:61)
    producer = None
    if streaming_file:
        # real CSV: the C++ producer thread tokenises, paces (-p) and fills pinned batches for this rank
        if not os.path.exists(cfg.training_data_file_path):
            raise FileNotFoundError(f"training data file not found: {cfg.training_data_file_path}")
        producer = native.StreamProducer(cfg.training_data_file_path, cfg.num_workers, rank * W, W, cfg.num_features,
                                         cfg.producer_time_per_event, batch_rows=eng.stage_rows, num_batches=4,
                                         packed="palette")  # bitmap + palette codes over PCIe when F % 32 == 0
        n_total, mine, cursor = 0, np.zeros(0, dtype=np.int64), 0
    else:
        n_total = len(train)
        gw = np.arange(n_total) % cfg.num_workers
        mine = np.nonzero((gw >= rank * W) & (gw < (rank + 1) * W))[0]
        cursor = int(resume_meta["rows_ingested"]) // max(world, 1) if resume_meta else 0
        dense = train.to_dense(np.float16)
        pin_rows = torch.zeros((eng.stage_rows, cfg.num_features), dtype=torch.float16).pin_memory()
        pin_lab = torch.zeros(eng.stage_rows, dtype=torch.int32).pin_memory()
    producer_eof = False
    delays = {int(a): float(b) for a, b in (x.split(":") for x in cfg.inject_delay.split(",") if x)}
    wd = Watchdog(lambda: eng.clock, eng.status, ns.watchdog_s).start()
    p_ms = cfg.producer_time_per_event
    t_start = time.time()
    warm = cfg.num_workers * 128                       # unthrottled warm-up rows (CsvProducer.java:73)
    eof_announced = False
    max_clock = cfg.max_clock
    try:
        while True:
            if producer is not None:
                # ---- drain whatever the producer thread has published
                for _ in range(4):
                    b, n, eof = producer.poll()
                    if b is None:
                        if eof == 2:
                            raise FileNotFoundError(cfg.training_data_file_path)
                        producer_eof = producer_eof or eof == 1
                        break
                    eng.ingest(producer.packed(b, n) if producer.is_packed else producer.rows[b, :n],
                               producer.labels[b, :n], producer.workers[b, :n].numpy(), producer.times[b, :n].numpy())
                    eng.ingest_done.synchronize()           # the pinned batch may be refilled after this
                    producer.release(b)
                    if p_ms <= 0:
                        break                               # unthrottled: one batch per clock
            # ---- in-memory slice: how many of my rows are due now?
            elif p_ms <= 0:
                due = min(len(mine), cursor + W * rows_per_clock)
            else:
                elapsed = time.time() - t_start
                rate = (1000.0 / p_ms) if p_ms <= 1000 else 1000.0 / p_ms
                global_due = min(n_total, warm + int(elapsed * rate))
                due = int(np.searchsorted(mine, global_due))
            if producer is None:
                due = min(due, cursor + eng.stage_rows)
            if producer is None and due > cursor:
                idx = mine[cursor:due]
                n = len(idx)
                pin_rows[:n].copy_(torch.from_numpy(dense[idx]))
                pin_lab[:n].copy_(torch.from_numpy(train.labels[idx].astype(np.int32)))
                now = np.full(n, (time.time() - t_start) * 1000.0) if p_ms > 0 else \
                    np.arange(n, dtype=np.float64) // max(W, 1)
                eng.ingest(pin_rows[:n], pin_lab[:n], (gw[idx] - rank * W).astype(np.int32), now)
                cursor = due
            done_feeding = producer_eof if producer is not None else cursor >= len(mine)
            if done_feeding and not eof_announced and rank == 0:
                server_log.println("CSV Producer sent all records!")
                eof_announced = True
            if eng.rows_ingested == 0:
                time.sleep(0.01)
                continue
            if rank in delays:
                time.sleep(delays[rank] / 1000.0)
            # ---- one clock
            clock = eng.clock
            eng.step()
            loss = eng.read_loss().numpy()
            ts = int(time.time() * 1000)
            if worker_log is not None:
                seen = [b.ring_total() for b in eng.buffers]
                for j in range(W):
                    f1w, accw = (-1.0, -1.0)
                    if ns.worker_metrics:                   # the reference scores every local fit on the test set
                        f1w, accw = eng.evaluate(eng.worker_local_weights(j))
                    worker_log.println(worker_row(ts, rank * W + j, clock, float(loss[j]), f1w, accw, seen[j]))
            if rank == 0 and clock % max(1, cfg.eval_every) == 0:
                f1, acc = eng.evaluate()
                server_log.println(server_row(int(time.time() * 1000), clock, f1, acc))
            if ns.checkpoint_dir and ns.checkpoint_every and rank == 0 and eng.clock % ns.checkpoint_every == 0 \
                    and cfg.consistency_model == 0:
                checkpoint.save(ns.checkpoint_dir, eng.clock, eng.weights().cpu().numpy(), layout,
                                rows_ingested=eng.rows_ingested * world)
            if wd.tripped:
                raise RuntimeError(wd.tripped)
            if max_clock >= 0 and eng.clock >= max_clock:
                break
            if max_clock < 0 and done_feeding:
                break
    finally:
        wd.stop()
        if producer is not None:
            producer.stop()
    if world > 1:
        dist.barrier()
    for s in (server_log, worker_log):
        if s is not None:
            s.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
