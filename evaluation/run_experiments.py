"""Re-runs the reference's eight evaluation experiments (evaluation/logs/*.csv of the reference:
single-worker_5tps, 4-workers_{0-5,2-5,5,10}tps, sequential, bounded_delay_10, eventual) on the
event-driven engine with a virtual clock, synthetic fine-food-shaped data and the CPU oracle math,
and writes logs with the same names and formats.  Deterministic; ~1 minute per run.

Worker iteration time is a model parameter of the virtual clock: 2.7 s (the median measured in the
reference's 4-workers_5tps log) plus a per-worker skew and a small deterministic jitter, which is what
makes the relaxed consistency models drift apart like they do in the reference (README.md:293-319).
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from psb200.config import PSConfig, SERVER_LOG_HEADER, WORKER_LOG_HEADER
from psb200.data import ReviewGenerator
from psb200.engine.sim import PSSimulation
from psb200.utils.logs import LogSink

RUNS = [  # name, workers, consistency, -p, virtual seconds
    ("single-worker_5tps", 1, 0, 200, 1053),
    ("4-workers_0-5tps", 4, 0, 2000, 1281),
    ("4-workers_2-5tps", 4, 0, 400, 1112),
    ("4-workers_5tps", 4, 0, 200, 517),
    ("4-workers_10tps", 4, 0, 100, 315),
    ("sequential", 4, 0, 200, 1946),
    ("bounded_delay_10", 4, 10, 200, 1872),
    ("eventual", 4, -1, 200, 1978),
]


def iter_time(worker, clock):
    jitter = ((worker * 7919 + clock * 104729) % 1000) / 1000.0 - 0.5
    return 2700.0 * (1.0 + 0.03 * worker) * (1.0 + 0.10 * jitter)


def main(only=None):
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "logs")
    gen = ReviewGenerator(mean_len=420.0)            # ~35 % dense rows, like the datawig embeddings
    train, test = gen.train_split(12000), gen.test_split()
    for name, workers, model, p, seconds in RUNS:
        if only and name not in only:
            continue
        t0 = time.time()
        cfg = PSConfig(num_workers=workers, consistency_model=model, producer_time_per_event=p)
        slog = LogSink.to_file(os.path.join(out, f"{name}_logs-server.csv"), SERVER_LOG_HEADER)
        wlog = LogSink.to_file(os.path.join(out, f"{name}_logs-worker.csv"), WORKER_LOG_HEADER)
        sim = PSSimulation(cfg, train, test, server_log=slog, worker_log=wlog, iter_time_ms=iter_time, virtual=True)
        sim.run(max_clock=10 ** 9, max_virtual_ms=seconds * 1000.0 + 21000.0)
        slog.close(), wlog.close()
        print(f"{name}: clocks={sim.server_clock} max_gap={sim.max_gap} rows={sim.rows_sent} ({time.time() - t0:.0f}s)",
              flush=True)


if __name__ == "__main__":
    main(sys.argv[1:] or None)
