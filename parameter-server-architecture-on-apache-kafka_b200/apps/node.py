"""Fused multi-GPU entry point: ``torchrun --nproc-per-node N -m psb200.apps.node [flags]``.

One process per GPU; every rank hosts a server shard, ``--workers_per_gpu`` workers and its slice of
the producer.  It accepts the union of the ServerAppRunner and WorkerAppRunner options
(``ServerAppRunner.java:19-26``, ``WorkerAppRunner.java:17-24``) with their defaults, writes the same
``logs-server.csv`` (rank 0) and ``logs-worker.csv`` (``logs-worker-<rank>.csv`` beyond rank 0) rows,
and runs the streaming loop on the device engine: producer slice -> pinned staging -> ring store ->
one fused step kernel per clock -> evaluation on rank 0.

The loop is the software pipeline ``bench.py`` times: the host->device copies of the next batches run
on the copy stream while a step trains (``ingest_begin`` / ``ingest_commit``), and the losses of clock
v are read on the host after clock v+1 has been queued (``read_loss_begin`` / ``read_loss_end``), so the
log rows of a clock are written one clock later.  Relaxed consistency models (``-c -1``, ``-c k``) run
with per-worker clocks: ``--clocks_per_launch`` clocks per launch, each worker at its own pace.

Extras over the reference: ``--checkpoint_dir/--checkpoint_every/--resume`` (all consistency models; ring
contents and buffer statistics included), ``--inject_delay worker:ms`` (straggler experiments, README 5.3),
``--watchdog_s``, ``--worker_metrics`` (test F1 / accuracy of every worker's local model, one launch).
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

SM_CYCLES_PER_MS = 1_900_000          # straggler injection in the per-worker modes sleeps on the SM clock


def build_parser():
    p = cli.server_parser()
    p.prog = "psb200.apps.node"
    p.add_argument("--workers_per_gpu", type=int, default=1)
    p.add_argument("--rows_per_clock", type=int, default=0,
                   help="rows each worker ingests per clock when -p 0 (default: min buffer size)")
    p.add_argument("--groups", type=int, default=1,
                   help="relaxed models: launch the workers of a GPU as this many independent groups (own streams)")
    p.add_argument("--checkpoint_dir", default="")
    p.add_argument("--checkpoint_every", type=int, default=0)
    p.add_argument("--resume", action="store_true")
    p.add_argument("--watchdog_s", type=float, default=120.0)
    p.add_argument("--worker_metrics", action="store_true",
                   help="score every worker's local model on the test set each clock (one batched launch)")
    return p


def main(argv: Optional[List[str]] = None) -> int:
    ns = cli.parse(build_parser(), argv)
    from ..utils.logging_config import configure
    configure()
    import torch
    import torch.distributed as dist
    from ..engine.device_engine import DeviceEngine
    from ..models.logreg import LogRegLayout
    from ..ops import native
    from ..utils import checkpoint
    from ..utils.watchdog import Watchdog

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from ..utils.numa import bind_to_device_node
    bind_to_device_node(local)                      # threads and pinned batches next to this rank's GPU (best effort)
    ctl = None                                      # host-side control group (stop agreement, barriers)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        ctl = dist.new_group(backend="gloo")
    W = ns.workers_per_gpu
    ns.num_workers = W * world
    cfg = cli.config_from_args(ns, "server")
    relaxed = cfg.consistency_model != 0
    cpl = max(1, ns.clocks_per_launch) if relaxed else 1
    if cfg.verbose and rank == 0:
        print("\nUsed parameter:")
        for k in ("training_data_file_path", "test_data_file_path", "consistency_model", "producer_time_per_event",
                  "min_buffer_size", "max_buffer_size", "buffer_size_coefficient", "num_workers"):
            print(f"    {k}: {getattr(cfg, k)}")
        print(f"    ranks: {world} x {W} workers, {cpl} clock(s) per launch\n")
    streaming_file = not cfg.training_data_file_path.startswith("synthetic:")
    train = None if streaming_file else load_csv_or_synthetic(cfg.training_data_file_path, cfg, "train")
    test = load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test")
    impl = native.IMPL_TC if ns.impl == "tc" else native.IMPL_SIMT
    if impl == native.IMPL_TC and (cfg.num_features % 128 or cfg.num_features > 1024 or cfg.max_buffer_size % 128):
        impl = native.IMPL_SIMT
    layout = LogRegLayout(cfg.num_features, cfg.num_classes)
    w0, resume_meta, resume_state = None, None, None
    if ns.resume and ns.checkpoint_dir:
        w0, resume_meta = checkpoint.load(ns.checkpoint_dir, layout)
        resume_state = checkpoint.load_rank_state(ns.checkpoint_dir, rank)
    elif cfg.init == "random":
        w0 = layout.random(cfg.seed)
    rows_per_clock = ns.rows_per_clock or cfg.min_buffer_size
    # a staging set never carries more rows for one worker than its ring holds
    stage_rows = min(max(W * rows_per_clock, 1024), W * cfg.max_buffer_size)
    eng = DeviceEngine(cfg, W, rank, world, impl, dev, w0=w0, stage_rows=stage_rows,
                       async_granularity="worker", groups=ns.groups if relaxed else 1,
                       max_clocks_per_launch=max(cpl, 1))
    eng.connect()
    if resume_state is not None:
        eng.load_state_dict(resume_state, weights=w0)
    elif resume_meta:                               # weights-only checkpoint: rings refill from the stream
        eng.clock = int(resume_meta["clock"])
        eng.reset_clock_words(eng.clock)
        eng.rows_ingested = int(resume_meta["rows_ingested"]) // max(world, 1)
    eng.load_test(test)
    server_log = worker_log = None
    if rank == 0:
        server_log = LogSink.to_file("./logs-server.csv", SERVER_LOG_HEADER) if cfg.logging else LogSink()
    wl_name = "./logs-worker.csv" if rank == 0 else f"./logs-worker-{rank}.csv"
    worker_log = LogSink.to_file(wl_name, WORKER_LOG_HEADER) if cfg.logging else (LogSink() if rank == 0 else None)
    if ns.metrics_port > 0:                         # one endpoint per rank: port + rank
        from ..utils.monitoring import start_exporter
        start_exporter(ns.metrics_port + rank, server_log, worker_log)

    # this rank's slice of the stream: global row i goes to worker i % numWorkers (CsvProducer.java:61)
    producer = None
    in_flight = []                                  # (producer batch, staging set) whose H2D may still be running
    if streaming_file:
        # real CSV: the C++ producer thread tokenises, paces (-p) and fills pinned batches for this rank
        if not os.path.exists(cfg.training_data_file_path):
            raise FileNotFoundError(f"training data file not found: {cfg.training_data_file_path}")
        producer = native.StreamProducer(cfg.training_data_file_path, cfg.num_workers, rank * W, W, cfg.num_features,
                                         cfg.producer_time_per_event, batch_rows=eng.stage_rows, num_batches=4,
                                         packed="palette",  # bitmap + palette codes over PCIe when F % 32 == 0
                                         has_header=cfg.csv_header)
        n_total, mine, cursor = 0, np.zeros(0, dtype=np.int64), 0
    else:
        n_total = len(train)
        gw = np.arange(n_total) % cfg.num_workers
        mine = np.nonzero((gw >= rank * W) & (gw < (rank + 1) * W))[0]
        cursor = int(resume_state["rows_ingested"]) if resume_state is not None else \
            (int(resume_meta["rows_ingested"]) // max(world, 1) if resume_meta else 0)
        dense = train.to_dense(np.float16)
        pins = [(torch.zeros((eng.stage_rows, cfg.num_features), dtype=torch.float16).pin_memory(),
                 torch.zeros(eng.stage_rows, dtype=torch.int32).pin_memory()) for _ in range(eng.stage_sets)]
        pin_i = 0
    producer_eof = False
    delays = {int(a): float(b) for a, b in (x.split(":") for x in cfg.inject_delay.split(",") if x)}
    if relaxed and delays:                          # per-worker modes: the straggler is a worker, delayed on the device
        wk, ms = next(iter(delays.items()))
        eng.inject_delay(wk, int(ms * SM_CYCLES_PER_MS))
    wd = Watchdog(lambda: eng.clock, eng.status, ns.watchdog_s).start()
    p_ms = cfg.producer_time_per_event
    t_start = time.time()
    warm = cfg.num_workers * 128                       # unthrottled warm-up rows (CsvProducer.java:73)
    eof_announced = False
    max_clock = cfg.max_clock
    pending = None                                  # (clock, loss ticket | None, numTuplesSeen) of the step just queued

    def release_landed(block: bool = False) -> None:
        """Hand pinned producer batches back once their host->device copies have landed."""
        while in_flight and (block or eng.h2d_done[in_flight[0][1]].query()):
            b, j = in_flight.pop(0)
            eng.h2d_done[j].synchronize()
            producer.release(b)

    def log_clock(rec) -> None:
        """Worker rows (and the server row) of a finished clock."""
        clock, ticket, seen, hist = rec
        ts = int(time.time() * 1000)
        if ticket is not None:
            losses = [(clock, eng.read_loss_end(ticket).numpy().copy())]
        else:                                       # per-worker mode: one record per clock of the launch and worker
            losses = [(int(hist[0]["clock"][c, 0]), np.concatenate([h["loss"][c] for h in hist]))
                      for c in range(hist[0].shape[0])]
        f1w = accw = None
        if ns.worker_metrics and worker_log is not None:
            f1w, accw = eng.evaluate_workers()      # the local models of the last clock, one launch
        if worker_log is not None:
            for ci, (c, loss) in enumerate(losses):
                last = ci == len(losses) - 1
                for j in range(W):
                    f1, acc = (float(f1w[j]), float(accw[j])) if (f1w is not None and last) else (-1.0, -1.0)
                    worker_log.println(worker_row(ts, rank * W + j, c, float(loss[j]), f1, acc, seen[j]))

    try:
        while True:
            # ---- feed: begin the host->device copies of whatever is due (up to the free staging sets)
            if producer is not None:
                release_landed()
                while eng.stage_pending < eng.stage_sets:
                    b, n, eof = producer.poll()
                    if b is None:
                        if eof == 2:
                            raise FileNotFoundError(cfg.training_data_file_path)
                        producer_eof = producer_eof or eof == 1
                        break
                    j = eng.stage_head
                    eng.ingest_begin(producer.packed(b, n) if producer.is_packed else producer.rows[b, :n],
                                     producer.labels[b, :n], producer.workers[b, :n].numpy(),
                                     producer.times[b, :n].numpy())
                    in_flight.append((b, j))
                    if p_ms <= 0:
                        break                               # unthrottled: one batch per clock
            else:
                if p_ms <= 0:
                    due = min(len(mine), cursor + W * rows_per_clock)
                else:
                    elapsed = time.time() - t_start
                    rate = (1000.0 / p_ms) if p_ms <= 1000 else 1000.0 / p_ms
                    global_due = min(n_total, warm + int(elapsed * rate))
                    due = int(np.searchsorted(mine, global_due))
                while due > cursor and eng.stage_pending < eng.stage_sets:
                    idx = mine[cursor:min(due, cursor + eng.stage_rows)]
                    n = len(idx)
                    pin_rows, pin_lab = pins[pin_i]
                    pin_i = (pin_i + 1) % len(pins)
                    pin_rows[:n].copy_(torch.from_numpy(dense[idx]))
                    pin_lab[:n].copy_(torch.from_numpy(train.labels[idx].astype(np.int32)))
                    now = np.full(n, (time.time() - t_start) * 1000.0) if p_ms > 0 else \
                        np.arange(n, dtype=np.float64) // max(W, 1)
                    eng.ingest_begin(pin_rows[:n], pin_lab[:n], (gw[idx] - rank * W).astype(np.int32), now)
                    cursor += n
            done_feeding = producer_eof if producer is not None else cursor >= len(mine)
            if done_feeding and not eof_announced and rank == 0:
                server_log.println("CSV Producer sent all records!")
                eof_announced = True
            # ---- every rank runs the same number of clocks: stop only when ALL ranks are done feeding, and
            #      wait (together) while some rank has not received a single row yet
            flags = torch.tensor([int(done_feeding), int(eng.rows_ingested > 0)], dtype=torch.int32)
            if ctl is not None:
                dist.all_reduce(flags, op=dist.ReduceOp.MIN, group=ctl)
            all_done, all_have_rows = bool(flags[0]), bool(flags[1])
            if not all_have_rows:
                if all_done:
                    break
                time.sleep(0.01)
                continue
            if not relaxed and rank in delays:
                time.sleep(delays[rank] / 1000.0)
            # ---- store the begun batches into the rings, then one clock (relaxed models: `cpl` clocks per worker)
            while eng.stage_pending:
                eng.ingest_commit()
            clock = eng.clock
            eng.step(num_clocks=cpl)
            seen = [b.ring_total() for b in eng.buffers]
            if relaxed:
                rec = (clock, None, seen, [eng.read_hist(g) for g in range(eng.groups)])   # synchronises
                if pending is not None:
                    log_clock(pending)
                pending = None
                log_clock(rec)
            else:
                rec = (clock, eng.read_loss_begin(), seen, None)
                if pending is not None:
                    log_clock(pending)              # clock v-1's rows, read while clock v runs
                pending = rec
            if rank == 0 and clock % max(1, cfg.eval_every) == 0:
                eng.wait_version()                  # all shards of this clock have reached the replica
                f1, acc = eng.evaluate()
                server_log.println(server_row(int(time.time() * 1000), eng.clock - 1, f1, acc))
            if ns.checkpoint_dir and ns.checkpoint_every and eng.clock % ns.checkpoint_every < cpl:
                eng.wait_version()
                torch.cuda.synchronize()
                if ctl is not None:
                    dist.barrier(group=ctl)         # every rank's pushes of this clock have landed
                checkpoint.save_rank_state(ns.checkpoint_dir, rank, eng.state_dict())
                if rank == 0:
                    checkpoint.save(ns.checkpoint_dir, eng.clock, eng.weights().cpu().numpy(), layout,
                                    rows_ingested=eng.rows_ingested * world,
                                    extra={"consistency_model": cfg.consistency_model, "world": world})
                if ctl is not None:
                    dist.barrier(group=ctl)
            if wd.tripped:
                raise RuntimeError(wd.tripped)
            if max_clock >= 0 and eng.clock >= max_clock:
                break
            if max_clock < 0 and all_done:
                break
        if pending is not None:
            log_clock(pending)
    finally:
        wd.stop()
        if producer is not None:
            try:
                release_landed(block=True)
            except Exception:
                pass
            producer.stop()
    if world > 1:
        dist.barrier(group=ctl)
    for s in (server_log, worker_log):
        if s is not None:
            s.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
