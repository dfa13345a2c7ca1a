"""Derived metrics of a set of evaluation logs, the way BASELINE.md derives them from the reference's:
duration = last - first timestamp; worker-its/s = worker rows / duration; F1 / accuracy from the server
log; gap = max difference between the workers' in-progress vector clocks when replaying the worker log.
Usage: python evaluation/summarize.py [logs_dir ...]   (default: evaluation/logs and, if present,
/root/reference/evaluation/logs) -> markdown table on stdout."""
import glob
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from psb200.utils.logs import parse_log


def load_log(prefix):
    return parse_log(prefix + "_logs-server.csv", "server"), parse_log(prefix + "_logs-worker.csv", "worker")


def max_clock_gap(worker_rows):
    rows = worker_rows[np.argsort(worker_rows[:, 0], kind="stable")]
    cur, gap = {}, 0
    for r in rows:
        cur[int(r[1])] = int(r[2])
        gap = max(gap, max(cur.values()) - min(cur.values()))
    return gap


def summarize(prefix):
    s, w = load_log(prefix)
    dur = (max(s[-1, 0], w[-1, 0]) - min(s[0, 0], w[0, 0])) / 1000.0
    workers = len(set(w[:, 1]))
    seen = {}
    for r in w:
        seen[int(r[1])] = r[6]
    return {"run": os.path.basename(prefix), "workers": workers, "server clocks": len(s), "duration s": round(dur),
            "worker-its/s": round(len(w) / dur, 2), "ingest tuples/s": round(sum(seen.values()) / dur, 2),
            "F1 final": round(s[-1, 4], 4), "F1 max": round(s[:, 4].max(), 4), "F1 mean last 50": round(s[-50:, 4].mean(), 4),
            "acc final": round(s[-1, 5], 4), "max clock gap": max_clock_gap(w)}


def table(directory):
    prefixes = sorted(p[:-len("_logs-server.csv")] for p in glob.glob(os.path.join(directory, "*_logs-server.csv")))
    rows = [summarize(p) for p in prefixes]
    if not rows:
        return ""
    cols = list(rows[0])
    out = ["| " + " | ".join(cols) + " |", "|" + "---|" * len(cols)]
    out += ["| " + " | ".join(str(r[c]) for c in cols) + " |" for r in rows]
    return "\n".join(out)


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    dirs = sys.argv[1:] or [os.path.join(here, "logs"), "/root/reference/evaluation/logs"]
    for d in dirs:
        if os.path.isdir(d):
            print(f"\n### {d}\n")
            print(table(d))
