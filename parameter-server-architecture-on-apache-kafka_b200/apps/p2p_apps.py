"""The two reference entry points on the product path: ``ServerAppRunner`` / ``WorkerAppRunner`` with
``--transport p2p`` (the default of ``--backend cuda``).

Process model of the reference (``run.sh:10-19``, ``ServerAppRunner.java:86-98``, ``WorkerApp.java:33-42``): a server
JVM (producer thread + parameter server) and a worker JVM (all ``numWorkers`` partitions), three Kafka topics in
between.  Here:

* the **server app** owns what the topics were: it allocates the peer-memory block (weights, mailboxes, clocks —
  ``csrc/common.cuh: CommLayout``) and the INPUT_DATA topic (``psb200.parallel.p2p.InputTopic*``) in device memory,
  publishes their CUDA IPC handles at the rendezvous address (``-r`` keeps its meaning), runs the producer
  (C++ tokenizer thread, ``-p`` pacing) whose packed batches it copies into topic slots, watches worker 0's clock in
  the block and writes one ``logs-server.csv`` row per clock with the test F1 / accuracy of the published weights
  (``ServerProcessor.java:154-165``);
* the **worker app** maps both, stores arriving batches into its input rings (dynamic ``-min/-max/-bc`` buffer,
  ``WorkerSamplingProcessor.java:50-135``) and runs the fused step kernel for all partitions: pull, local solve,
  push and apply happen inside the kernel, in the server's memory (``WorkerTrainingProcessor.java:63-98``,
  ``ServerProcessor.java:143-183``).  It writes the ``logs-worker.csv`` rows.

Rows, weights and gradients never pass through a host socket; the rendezvous is used once, for the handles and a
JSON description of the run (the constants both JVMs share in the reference, ``BaseKafkaApp.java:25-36``).
"""
from __future__ import annotations

import threading
import time
from typing import Optional

import numpy as np

from ..config import PSConfig
from ..utils.logs import LogSink, server_row, worker_row

RUN_KEY, COMM_KEY, INPUT_KEY, READY_KEY = "psb/run", "psb/comm_handle", "psb/input_handle", "psb/worker_ready"


# ------------------------------------------------------------------ producers (server side)
class MemoryProducer:
    """``StreamProducer`` interface over an in-memory data set (``synthetic:<n>`` paths): round-robin partitioning
    (``CsvProducer.java:61``), the reference's pacing (``:73-83``: numWorkers*128 rows unthrottled, then 1000/p rows
    per second in total; p = 0 never sleeps), palette-packed batches in pinned memory."""

    def __init__(self, train, num_workers: int, num_features: int, p_ms: int, batch_rows: int, num_batches: int = 4):
        from ..ops import native
        self.dense = train.to_dense(np.float16)
        self.labels = train.labels.astype(np.int32)
        self.W, self.p_ms, self.batch_rows = num_workers, p_ms, batch_rows
        import torch
        self.slots = [native.PackedRows(batch_rows, num_features, palette=True) for _ in range(num_batches)]
        self.meta = [(torch.zeros(batch_rows, dtype=torch.int32).pin_memory(),
                      torch.zeros(batch_rows, dtype=torch.float64).pin_memory()) for _ in range(num_batches)]
        self.free = list(range(num_batches))
        self.cursor, self.t0 = 0, time.time()
        self.is_packed = True

    def _due(self) -> int:
        n = len(self.labels)
        if self.p_ms <= 0:
            return min(n, self.cursor + self.batch_rows)
        warm = self.W * 128
        rate = 1000.0 / self.p_ms
        return min(n, warm + int((time.time() - self.t0) * rate), self.cursor + self.batch_rows)

    def poll(self):
        if self.cursor >= len(self.labels):
            return None, 0, 1
        due = self._due()
        if due <= self.cursor or not self.free:
            return None, 0, 0
        b = self.free.pop(0)
        idx = np.arange(self.cursor, due)
        self.slots[b].pack_dense(self.dense[idx], self.labels[idx])
        now = (time.time() - self.t0) * 1000.0
        times = np.full(len(idx), now) if self.p_ms > 0 else (idx // max(self.W, 1)).astype(np.float64)
        workers, tms = self.meta[b]
        workers.numpy()[: len(idx)] = idx % self.W
        tms.numpy()[: len(idx)] = times
        self.cursor = due
        return b, len(idx), 0

    def batch(self, b: int, n: int):
        pk = self.slots[b]
        workers, times = self.meta[b]
        return n, workers, times, pk.labels, pk.bitmap, pk.row_off, pk.payload, pk.payload_bytes

    def release(self, b: int) -> None:
        self.free.append(b)

    def stop(self) -> None:
        pass


class FileProducer:
    """The C++ producer thread (``psb_producer_start_palette``) behind the same small interface."""

    def __init__(self, path: str, num_workers: int, num_features: int, p_ms: int, batch_rows: int, num_batches: int = 4,
                 has_header: bool = True):
        from ..ops import native
        self.p = native.StreamProducer(path, num_workers, 0, num_workers, num_features, p_ms, batch_rows=batch_rows,
                                       num_batches=num_batches, packed="palette", has_header=has_header)

    def poll(self):
        return self.p.poll()

    def batch(self, b: int, n: int):
        p = self.p
        return (n, p.workers[b], p.times[b], p.labels[b], p.bitmap[b], p.row_off[b], p.payload[b],
                int(p.row_off[b, n]))

    def release(self, b: int) -> None:
        self.p.release(b)

    def stop(self) -> None:
        self.p.stop()


# ------------------------------------------------------------------ server app
class P2PServer:
    def __init__(self, cfg: PSConfig, server_log: LogSink, device: int = 0, address: Optional[str] = None,
                 clocks_per_launch: int = 1, impl: str = "tc", worker_metrics: bool = True, server_sync: bool = True):
        import torch
        from ..models.logreg import LogRegLayout
        from ..ops import native
        from ..parallel import p2p
        from .base_app import load_csv_or_synthetic
        self.cfg, self.log, self.torch, self.native, self.p2p = cfg, server_log, torch, native, p2p
        torch.cuda.set_device(device)
        self.dev = torch.device("cuda", device)
        self.layout = LogRegLayout(cfg.num_features, cfg.num_classes)
        F, W = cfg.num_features, cfg.num_workers
        self.F, self.nv = F, self.layout.device_size
        w0 = self.layout.zeros() if cfg.init == "zero" else self.layout.random(cfg.seed)
        self.w0 = torch.from_numpy(w0.reshape(-1).copy()).to(self.dev)
        # "createTopics" (ServerApp.java:31-42): the block and the input topic
        self.block_bytes = native.comm_block_bytes(F, 1)
        self.block = native.device_malloc(self.block_bytes)
        native.check(native.lib().psb_comm_init_block(self.block, self.w0.data_ptr(), F, 1, 0, native._stream()),
                     "psb_comm_init_block")
        rows = max(1024, W * 128)
        self.topic_layout = p2p.InputTopicLayout(F, rows, 4)
        self.region = p2p.DeviceRegion(self.topic_layout.bytes)
        self.writer = p2p.InputTopicWriter(self.region, self.topic_layout)
        # evaluation of the published weights
        test = load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test")
        self.test_X = torch.from_numpy(test.to_dense(np.float16)).to(self.dev)
        self.test_y = torch.from_numpy(test.labels.astype(np.int32)).to(self.dev)
        self.snap = torch.zeros(self.nv, dtype=torch.float32, device=self.dev)
        self.conf = torch.zeros(64, dtype=torch.int32, device=self.dev)
        self.metrics = torch.zeros(2, dtype=torch.float32, device=self.dev)
        self.words = torch.zeros(16, dtype=torch.int32).pin_memory()
        self.eval_impl = native.IMPL_TC if F % 64 == 0 else native.IMPL_SIMT
        L = native.lib()
        self.p_version = L.psb_comm_version_words(self.block, F, 1)
        self.p_wclock = L.psb_comm_worker_clocks(self.block, F, 1)
        self.p_status = L.psb_comm_status(self.block, F, 1)
        self.p_pub, self.p_pub_seq = L.psb_comm_published(self.block, F, 1), L.psb_comm_published_seq(self.block, F, 1)
        self.p_auth = L.psb_comm_shard_weights(self.block, F, 1)
        self.address = address or cfg.rendezvous
        self.description = {
            "num_workers": W, "num_features": F, "num_classes": cfg.num_classes,
            "consistency_model": cfg.consistency_model, "local_solver": cfg.local_solver, "max_iter": cfg.max_iter,
            "standardize": cfg.standardize, "center": cfg.center, "sgd_lr": cfg.sgd_lr, "max_clock": cfg.max_clock,
            "inject_delay": cfg.inject_delay, "clocks_per_launch": clocks_per_launch, "impl": impl,
            "worker_metrics": worker_metrics, "server_sync": server_sync, "server_device": device, "block_bytes": self.block_bytes,
            "topic": self.topic_layout.describe(), "topic_bytes": self.topic_layout.bytes,
        }
        self.closing = False
        self.watchdog_s = 120.0
        torch.cuda.synchronize()

    # -- transport set-up
    def announce(self):
        """Publish the run description and the IPC handles at the broker address."""
        rv = self.p2p.Rendezvous(self.address, is_server=True)
        rv.put(RUN_KEY, self.description)
        rv.put(COMM_KEY, self.native.ipc_export(self.block))
        rv.put(INPUT_KEY, self.region.handle())
        return rv

    def _read_words(self, ptr: int, n: int) -> np.ndarray:
        st = self.torch.cuda.current_stream()
        self.native.check(self.native.lib().psb_d2h_copy(self.words.data_ptr(), ptr, 4 * n, st.cuda_stream), "psb_d2h_copy")
        st.synchronize()
        return self.words[:n].numpy().copy()

    def _snapshot(self):
        """(clocks applied, consistent?) and the weights in self.snap.  Sequential: the seqlock-published snapshot
        (retried when a publish was in progress); relaxed: the live shard, as it is (ServerProcessor.java:172-182)."""
        L, st = self.native.lib(), self.torch.cuda.current_stream().cuda_stream
        if self.cfg.consistency_model == 0:
            s1 = int(self._read_words(self.p_pub_seq, 1)[0])
            self.native.check(L.psb_d2d_copy(self.snap.data_ptr(), self.p_pub, self.nv * 4, st), "psb_d2d_copy")
            s2 = int(self._read_words(self.p_pub_seq, 1)[0])
            return s1 // 2, (s1 == s2 and s1 % 2 == 0)
        clock = int(self._read_words(self.p_wclock, 1)[0])
        self.native.check(L.psb_d2d_copy(self.snap.data_ptr(), self.p_auth, self.nv * 4, st), "psb_d2d_copy")
        return clock, True

    def _log_row(self, clock_done: int) -> None:
        self.native.evaluate(self.test_X, self.test_y, self.test_X.shape[0], self.snap, self.F, self.layout.R, self.conf,
                             self.metrics, self.eval_impl)
        f1, acc = self.metrics.cpu().tolist()
        self.log.println(server_row(int(time.time() * 1000), clock_done - 1, f1, acc))
        self.writer.set_flag(self.p2p.CTL_SERVER_LOGGED, clock_done)

    def run(self, producer) -> None:
        """Feed the topic and log until the worker app reports its last clock."""
        w, cfg = self.writer, self.cfg
        held = None                                      # a polled batch waiting for a free slot
        in_flight = []                                   # (producer batch, event) copies still reading pinned memory
        eof_sent, logged = False, 0
        seen_state, seen_at = None, time.time()          # watchdog: what the worker app last reported, and when
        try:
            while True:
                progressed = False
                # ---- INPUT_DATA: producer batches -> topic slots
                while in_flight and in_flight[0][1].query():
                    producer.release(in_flight.pop(0)[0])
                if not eof_sent:
                    if held is None:
                        b, n, eof = producer.poll()
                        if b is not None:
                            held = (b, n)
                        elif eof == 2:
                            raise FileNotFoundError(cfg.training_data_file_path)
                        elif eof == 1:
                            w.set_flag(self.p2p.CTL_EOF)
                            self.log.println("CSV Producer sent all records!")
                            eof_sent = True
                    if held is not None and w.has_room():
                        b, n = held
                        assert w.publish(*producer.batch(b, n))
                        ev = self.torch.cuda.Event()
                        ev.record(self.region.stream)
                        in_flight.append((b, ev))
                        held, progressed = None, True
                # ---- server log: one row per clock of worker 0 that we get to see
                clock_done, consistent = self._snapshot()
                if consistent and clock_done > logged:
                    self._log_row(clock_done)
                    logged, progressed = clock_done, True
                status = int(self._read_words(self.p_status, 1)[0])
                if status != 0:
                    raise RuntimeError(f"a worker kernel gave up waiting (status word {status})")
                consumed, worker_done, worker_clock = w.worker_state()
                if (consumed, worker_clock) != seen_state:
                    seen_state, seen_at = (consumed, worker_clock), time.time()
                elif self.watchdog_s > 0 and time.time() - seen_at > self.watchdog_s and (w.produced > consumed or logged):
                    raise RuntimeError(f"the worker app made no progress for {self.watchdog_s:.0f}s "
                                       f"(batches consumed {consumed} of {w.produced}, clock {worker_clock})")
                if worker_done:
                    clock_done, consistent = self._snapshot()
                    if consistent and clock_done > logged:
                        self._log_row(clock_done)
                    break
                if self.closing:
                    w.set_flag(self.p2p.CTL_STOP)
                if not progressed:
                    time.sleep(0.0005)
        finally:
            w.set_flag(self.p2p.CTL_STOP)
            self.torch.cuda.synchronize()

    def close(self) -> None:
        self.torch.cuda.synchronize()
        self.region.close()
        if self.block:
            self.native.device_free(self.block)
            self.block = 0


# ------------------------------------------------------------------ worker app
class P2PWorker:
    def __init__(self, wcfg: PSConfig, worker_log: LogSink, device: int = 0, address: Optional[str] = None,
                 local: Optional[dict] = None):
        """wcfg: the WorkerAppRunner options (-test, -min, -max, -bc); everything both apps must agree on comes from
        the server app's run description.  local: {"run": description, "block": ptr, "region": ptr} when the worker is
        hosted by the server process (--in_process) — the same memory without IPC."""
        import torch
        from ..engine.device_engine import DeviceEngine
        from ..ops import native
        from ..parallel import p2p
        from .base_app import load_csv_or_synthetic
        self.torch, self.native, self.p2p, self.log = torch, native, p2p, worker_log
        torch.cuda.set_device(device)
        dev = torch.device("cuda", device)
        self.rv = None
        if local is None:
            self.rv = p2p.Rendezvous(address or wcfg.rendezvous, is_server=False)
            self.rv.wait_for(INPUT_KEY)
            run = self.rv.get_json(RUN_KEY)
            block = native.ipc_import(self.rv.get_bytes(COMM_KEY))
            self.region = p2p.DeviceRegion.from_handle(self.rv.get_bytes(INPUT_KEY), run["topic_bytes"])
            self._imported_block = block
        else:
            run, block = local["run"], local["block"]
            self.region = p2p.DeviceRegion(run["topic_bytes"], ptr=local["region"])
            self._imported_block = 0
        self.run_desc = run
        cfg = wcfg.replace(num_workers=run["num_workers"], num_features=run["num_features"],
                           num_classes=run["num_classes"], consistency_model=run["consistency_model"],
                           local_solver=run["local_solver"], max_iter=run["max_iter"], standardize=run["standardize"],
                           center=run["center"], sgd_lr=run["sgd_lr"], max_clock=run["max_clock"],
                           inject_delay=run["inject_delay"])
        self.cfg = cfg
        F, W = cfg.num_features, cfg.num_workers
        impl = native.IMPL_TC if run["impl"] == "tc" else native.IMPL_SIMT
        if impl == native.IMPL_TC and (F % 128 or F > 1024 or cfg.max_buffer_size % 128):
            impl = native.IMPL_SIMT
        self.cpl = max(1, int(run["clocks_per_launch"])) if cfg.consistency_model != 0 else 1
        tl = run["topic"]
        self.topic_layout = p2p.InputTopicLayout(tl["num_features"], tl["rows_per_set"], tl["num_sets"])
        self.reader = p2p.InputTopicReader(self.region, self.topic_layout)
        self.eng = DeviceEngine(cfg, W, 0, 1, impl, dev, stage_rows=tl["rows_per_set"], async_granularity="worker",
                                max_clocks_per_launch=self.cpl, external_block=block,
                                home_remote=run["server_device"] != device)
        self.eng.publish = True
        self.eng.load_test(load_csv_or_synthetic(cfg.test_data_file_path, cfg, "test"))
        self.worker_metrics = bool(run["worker_metrics"])
        self.server_sync = bool(run["server_sync"])
        delays = {int(a): float(b) for a, b in (x.split(":") for x in cfg.inject_delay.split(",") if x)}
        self.host_delay_s = 0.0
        if delays:
            wk, ms = next(iter(delays.items()))
            if cfg.consistency_model != 0:
                self.eng.inject_delay(wk, int(ms * 1_900_000))
            else:
                self.host_delay_s = ms / 1000.0
        self.closing = False
        if self.rv is not None:
            self.rv.put(READY_KEY, "1")

    def _take_batches(self) -> int:
        n = 0
        while True:
            b = self.reader.poll()
            if b is None:
                return n
            s = self.eng.ingest_from_device(b.n, b.workers, b.times, b.bitmap, b.row_off, b.payload, b.labels)
            self.reader.mark_consumed(b.seq, stream=s)           # behind the store kernel: the slot may be refilled
            n += 1

    def _log(self, clock0: int, losses, seen, f1w, accw) -> None:
        ts = int(time.time() * 1000)
        for ci, loss in enumerate(losses):
            last = ci == len(losses) - 1
            for j in range(self.eng.W):
                f1, acc = (float(f1w[j]), float(accw[j])) if (f1w is not None and last) else (-1.0, -1.0)
                self.log.println(worker_row(ts, j, clock0 + ci, float(loss[j]), f1, acc, seen[j]))

    def run(self) -> None:
        eng, cfg, reader = self.eng, self.cfg, self.reader
        relaxed = cfg.consistency_model != 0
        ran_after_last_batch = False
        try:
            while not self.closing:
                took = self._take_batches()
                if reader.stop:
                    break
                if eng.rows_ingested == 0:
                    if reader.drained:
                        break
                    time.sleep(0.001)
                    continue
                if took:
                    ran_after_last_batch = False
                elif cfg.max_clock < 0 and reader.drained and ran_after_last_batch:
                    break                                     # the stream is over and its last rows were trained on
                if self.host_delay_s:
                    time.sleep(self.host_delay_s)
                if self.server_sync:                          # reference: the server scores the weights of clock v
                    t_wait = time.time()
                    while not self.closing:                   # before it hands them out (ServerProcessor.java:154-182)
                        ctl = self.region.read(0, 16).view(np.int32)
                        if ctl[self.p2p.CTL_SERVER_LOGGED] >= eng.clock or ctl[self.p2p.CTL_STOP]:
                            break
                        if time.time() - t_wait > 120.0:
                            raise RuntimeError(f"the server app has not logged clock {eng.clock - 1} for 120 s")
                        time.sleep(0.0002)
                clock0 = eng.clock
                eng.step(num_clocks=self.cpl)
                seen = [b.ring_total() for b in eng.buffers]
                if relaxed:
                    h = eng.read_hist()
                    losses = [h["loss"][c] for c in range(h.shape[0])]
                else:
                    losses = [eng.read_loss().numpy().copy()]
                f1w = accw = None
                if self.worker_metrics:                       # the reference scores every local fit on the test set
                    f1w, accw = eng.evaluate_workers()
                self._log(clock0, losses, seen, f1w, accw)
                reader.set_flag(self.p2p.CTL_WORKER_CLOCK, eng.clock)
                ran_after_last_batch = True
                if cfg.max_clock >= 0 and eng.clock >= cfg.max_clock:
                    break
        finally:
            self.torch.cuda.synchronize()
            reader.set_flag(self.p2p.CTL_WORKER_DONE, 1)
            self.region.stream.synchronize()

    def close(self) -> None:
        self.eng.close()
        self.region.close()
        if self._imported_block:
            self.native.lib().psb_ipc_close(self._imported_block)
            self._imported_block = 0


# ------------------------------------------------------------------ entry points used by the runners
def make_producer(cfg: PSConfig, batch_rows: int):
    import os
    from .base_app import load_csv_or_synthetic
    path = cfg.training_data_file_path
    if path.startswith("synthetic:"):
        return MemoryProducer(load_csv_or_synthetic(path, cfg, "train"), cfg.num_workers, cfg.num_features,
                              cfg.producer_time_per_event, batch_rows)
    if not os.path.exists(path):
        raise FileNotFoundError(f"training data file not found: {path}")
    return FileProducer(path, cfg.num_workers, cfg.num_features, cfg.producer_time_per_event, batch_rows,
                        has_header=cfg.csv_header)


def run_server(app, cfg: PSConfig, ns, server_log: LogSink, worker_log: Optional[LogSink]) -> None:
    """ServerAppRunner --transport p2p.  With --in_process the worker app runs on a thread of this process."""
    srv = P2PServer(cfg, server_log, device=ns.device, address=ns.rendezvous or cfg.rendezvous,
                    clocks_per_launch=ns.clocks_per_launch, impl=ns.impl, worker_metrics=not ns.no_worker_metrics,
                    server_sync=not ns.free_running)
    app.on_close = lambda: setattr(srv, "closing", True)
    producer = make_producer(cfg, srv.topic_layout.rows_per_set)
    worker_thread, worker, failure = None, None, []
    try:
        if ns.in_process:
            worker = P2PWorker(cfg, worker_log or LogSink(), device=ns.device,
                               local={"run": srv.description, "block": srv.block, "region": srv.region.ptr})

            def _worker_main():
                try:
                    srv.torch.cuda.set_device(ns.device)
                    worker.run()
                except BaseException as e:                   # surfaced by the server loop
                    failure.append(e)
                    worker.reader.set_flag(srv.p2p.CTL_WORKER_DONE, 1)
            worker_thread = threading.Thread(target=_worker_main, name="psb-worker-app", daemon=True)
            worker_thread.start()
        else:
            rv = srv.announce()
            rv.wait_for(READY_KEY, timeout_s=600.0)
        if ns.start_delay_ms > 0:
            time.sleep(ns.start_delay_ms / 1000.0)
        srv.run(producer)
        if worker_thread is not None:
            worker_thread.join(timeout=60)
        if failure:
            raise failure[0]
    finally:
        producer.stop()
        if worker is not None:
            worker.close()
        srv.close()


def run_worker(app, wcfg: PSConfig, ns, worker_log: LogSink) -> None:
    """WorkerAppRunner --transport p2p."""
    wk = P2PWorker(wcfg, worker_log, device=ns.device, address=ns.rendezvous or wcfg.rendezvous)
    app.on_close = lambda: setattr(wk, "closing", True)
    try:
        wk.run()
    finally:
        wk.close()
