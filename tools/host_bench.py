"""Host-side throughput of the C++ runtime (no GPU needed): CSV tokenizer, row packers, slot assignment.

These are the host stages of ingestion (SURVEY hot spots H8 / H9: CsvProducer.java:46-61 parse + JSON,
WorkerSamplingProcessor.java:50-113 buffer maintenance).  One JSON line per stage; `--rows` rows of the review shape.

    python tools/host_bench.py [--rows 8192] [--reps 5]
"""
import argparse
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from psb200.data import ReviewGenerator, write_csv
from psb200.ops import native


def best(fn, reps):
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    return min(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=8192)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--workers", type=int, default=296)
    a = ap.parse_args()
    gen = ReviewGenerator(seed=1)
    batch = gen.train_split(a.rows)
    F = batch.num_features
    dense = np.ascontiguousarray(batch.to_dense(np.float16))
    out = []

    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "rows.csv")
        write_csv(path, batch)
        size = os.path.getsize(path)
        t = best(lambda: native.csv_to_csr_two_pass(path), a.reps)
        out.append({"stage": "csv_count+parse (two passes)", "rows_per_s": a.rows / t, "MB_per_s": size / t / 1e6,
                    "file_MB": size / 1e6})
        for threads in (1, 4, 8):
            t = best(lambda: native.csv_to_csr(path, threads=threads), a.reps)
            out.append({"stage": f"csv_open/fill (one pass, mapped file, {threads} thread{'s' if threads > 1 else ''})",
                        "rows_per_s": a.rows / t, "MB_per_s": size / t / 1e6})
        for packed, parts, threads in ((False, 1, 1), (False, 8, 1), ("palette", 1, 1), ("palette", 1, 2),
                                       ("palette", 1, 4)):
            # parts: this rank keeps every row / one partition in eight; threads: PSB_PRODUCER_THREADS
            def stream():
                os.environ["PSB_PRODUCER_THREADS"] = str(threads)
                prod = native.StreamProducer(path, parts, 0, 1, F, 0, batch_rows=512, num_batches=4, packed=packed)
                got = 0
                while True:
                    b, n, eof = prod.poll()
                    if b is None:
                        if eof:
                            break
                        time.sleep(0.0002)
                        continue
                    got += n
                    prod.release(b)
                prod.stop()
                assert got == (a.rows + parts - 1) // parts, got
            t = best(stream, max(3, a.reps))
            out.append({"stage": f"producer ({'palette-packed' if packed else 'dense'} batches, keeps 1 row in "
                                 f"{parts}, {threads} tokenising thread{'s' if threads > 1 else ''})",
                        "file_rows_per_s": a.rows / t, "kept_rows_per_s": a.rows / parts / t})
        os.environ.pop("PSB_PRODUCER_THREADS", None)
        # the same rows converted once into a rows file (tools/csv_to_rows.py): nothing to tokenise
        from psb200.data import write_rows_file
        rpath = os.path.join(d, "rows.psbrows")
        rsize = write_rows_file(rpath, batch)
        for parts in (1, 8):
            def stream_rows():
                prod = native.StreamProducer(rpath, parts, 0, 1, F, 0, batch_rows=512, num_batches=4, packed="palette")
                got = 0
                while True:
                    b, n, eof = prod.poll()
                    if b is None:
                        if eof:
                            break
                        time.sleep(0.0002)
                        continue
                    got += n
                    prod.release(b)
                prod.stop()
                assert got == (a.rows + parts - 1) // parts, got
            t = best(stream_rows, max(3, a.reps))
            out.append({"stage": f"producer, rows file (palette-packed batches, keeps 1 row in {parts})",
                        "file_rows_per_s": a.rows / t, "kept_rows_per_s": a.rows / parts / t,
                        "bytes_per_row": rsize / a.rows})

    for palette in (False, True):
        pk = native.PackedRows(a.rows, F, pin=False, palette=palette)
        t = best(lambda: pk.pack_dense(dense, batch.labels), a.reps)
        out.append({"stage": "pack_palette_dense" if palette else "pack_bitmap_dense", "rows_per_s": a.rows / t,
                    "in_GB_per_s": dense.nbytes / t / 1e9, "bytes_per_row": pk.nbytes / a.rows})
    pk = native.PackedRows(a.rows, F, pin=False)
    t = best(lambda: pk.pack_csr(batch), a.reps)
    out.append({"stage": "pack_bitmap_csr", "rows_per_s": a.rows / t})

    W = a.workers
    bufs = [native.NativeBuffer(128, 1024, 0.3, ring=1024 + 64) for _ in range(W)]
    handles = native.buffer_handles(bufs)
    n = W * 16
    workers = np.repeat(np.arange(W, dtype=np.int32), 16)
    ws = np.zeros((n, 2), dtype=np.int32)
    win = np.zeros((W, 2), dtype=np.int32)
    ch = np.zeros((W, 4), dtype=np.int32)
    clock = [0.0]

    def slots(order):
        clock[0] += 1000.0
        now = clock[0] + np.arange(n, dtype=np.float64) * 0.01
        native.buffers_on_batch(bufs, order, now, ws, win, handles=handles, change=ch)

    t = best(lambda: slots(workers), max(a.reps, 20))
    out.append({"stage": "buffers_on_batch (grouped)", "events_per_s": n / t, "us_per_batch": t * 1e6, "events": n})
    rr = np.tile(np.arange(W, dtype=np.int32), 16)
    t = best(lambda: slots(rr), max(a.reps, 20))
    out.append({"stage": "buffers_on_batch (round robin)", "events_per_s": n / t, "us_per_batch": t * 1e6, "events": n})

    for o in out:
        print(json.dumps({k: (round(v, 3) if isinstance(v, float) else v) for k, v in o.items()}))


if __name__ == "__main__":
    main()
