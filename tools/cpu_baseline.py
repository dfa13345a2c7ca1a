"""Timed CPU baseline: the reference's algorithm restated on the CPU (NumPy/SciPy oracle), same step as bench.py
(every worker: 2-iteration L-BFGS on its 1024-row buffer; w += sum/N).  Prints one JSON line.  The JVM/Kafka/Spark
reference itself cannot run here; BASELINE.md's ~130 samples/s comes from its published logs."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from psb200.data import ReviewGenerator
from psb200.models import LogRegLayout
from psb200.oracle import LogisticRegressionTask

workers = int(sys.argv[1]) if len(sys.argv) > 1 else 4
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
clocks = int(sys.argv[3]) if len(sys.argv) > 3 else 5
gen = ReviewGenerator(mean_len=420.0)
data = gen.train_split(workers * rows)
bufs = [data.slice(np.arange(w * rows, (w + 1) * rows)) for w in range(workers)]
lay = LogRegLayout()
W = lay.random(0)
tasks = [LogisticRegressionTask() for _ in range(workers)]
t0 = time.time()
for _ in range(clocks):
    tot = np.zeros_like(W, dtype=np.float64)
    for t, b in zip(tasks, bufs):
        t.set_weights(W)
        tot += t.calculate_gradients(b, eval_test=False)
    W = (W + tot / workers).astype(np.float32)
dt = time.time() - t0
print(json.dumps({"impl": "cpu-oracle (NumPy/SciPy restatement of the reference algorithm)", "workers": workers,
                  "buffer_rows": rows, "clocks": clocks, "samples_per_sec": workers * rows * clocks / dt,
                  "ms_per_clock": dt / clocks * 1e3, "cpu_count": os.cpu_count()}))
