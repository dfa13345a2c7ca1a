"""Plots of the evaluation logs (loss / weighted F1 / accuracy against tuples seen), headless.

Counterpart of the reference's plot-generation.ipynb / evaluation-multipleDatasetsAtOnce.ipynb: server
and worker logs are merged on the vector clock, the x axis is the sum of the workers' numTuplesSeen at
that clock.  matplotlib is used when importable; otherwise a small built-in rasteriser writes the PNGs
(no graphics stack is needed either way).  Usage: python evaluation/plot_generation.py [logs_dir [out_dir]]
"""
import glob
import os
import struct
import sys
import zlib

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from psb200.utils.logs import parse_log

COLORS = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189), (140, 86, 75),
          (227, 119, 194), (127, 127, 127)]


def load_run(prefix):
    s = parse_log(prefix + "_logs-server.csv", "server")
    w = parse_log(prefix + "_logs-worker.csv", "worker")
    clocks = s[:, 2].astype(int)
    seen, loss = [], []
    for c in clocks:
        rows = w[w[:, 2] == c]
        seen.append(rows[:, 6].sum() if len(rows) else np.nan)
        loss.append(rows[:, 3].mean() if len(rows) else np.nan)
    return {"x": np.asarray(seen), "f1": s[:, 4], "accuracy": s[:, 5], "loss": np.asarray(loss)}


def write_png(path, img):
    h, w, _ = img.shape
    raw = b"".join(b"\x00" + img[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def raster_plot(path, series, width=900, height=540):
    """series: list of (label, x, y).  Axes box, tick marks, one polyline per series."""
    img = np.full((height, width, 3), 255, dtype=np.uint8)
    l, r, t, b = 60, width - 20, 20, height - 40
    img[t:b, l] = img[t:b, r] = 0
    img[t, l:r] = img[b, l:r] = 0
    xs = np.concatenate([s[1][np.isfinite(s[1])] for s in series])
    ys = np.concatenate([s[2][np.isfinite(s[2])] for s in series])
    x0, x1, y0, y1 = xs.min(), xs.max(), ys.min(), ys.max()
    x1, y1 = (x1 if x1 > x0 else x0 + 1), (y1 if y1 > y0 else y0 + 1)
    for k in range(1, 5):                                   # light grid
        img[t + (b - t) * k // 5, l:r:4] = 200
        img[t:b:4, l + (r - l) * k // 5] = 200
    for i, (_, x, y) in enumerate(series):
        ok = np.isfinite(x) & np.isfinite(y)
        px = (l + (x[ok] - x0) / (x1 - x0) * (r - l)).astype(int)
        py = (b - (y[ok] - y0) / (y1 - y0) * (b - t)).astype(int)
        col = COLORS[i % len(COLORS)]
        for j in range(len(px) - 1):
            n = max(abs(px[j + 1] - px[j]), abs(py[j + 1] - py[j]), 1)
            lx = np.linspace(px[j], px[j + 1], n + 1).astype(int).clip(l, r)
            ly = np.linspace(py[j], py[j + 1], n + 1).astype(int).clip(t, b)
            img[ly, lx] = col
            img[(ly + 1).clip(t, b), lx] = col
        img[t + 6 + 10 * i: t + 12 + 10 * i, r - 40: r - 10] = col     # legend swatch, order = series order
    write_png(path, img)


def plot(metric, runs, path):
    series = [(name, d["x"], d[metric]) for name, d in runs.items()]
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        plt.figure(figsize=(9, 5.4))
        for name, x, y in series:
            plt.plot(x, y, label=name)
        plt.xlabel("tuples seen (sum over workers)"), plt.ylabel(metric), plt.legend(), plt.grid(alpha=0.3)
        plt.savefig(path, dpi=100), plt.close()
    except ImportError:
        raster_plot(path, series)


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    logs = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, "logs")
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(here, "plots")
    os.makedirs(out, exist_ok=True)
    prefixes = sorted(p[:-len("_logs-server.csv")] for p in glob.glob(os.path.join(logs, "*_logs-server.csv")))
    runs = {os.path.basename(p): load_run(p) for p in prefixes}
    groups = {"consistency": ["sequential", "bounded_delay_10", "eventual"],
              "event_rate": ["4-workers_0-5tps", "4-workers_2-5tps", "4-workers_5tps", "4-workers_10tps"],
              "num_workers": ["single-worker_5tps", "4-workers_5tps"]}
    for gname, names in groups.items():
        sel = {n: runs[n] for n in names if n in runs}
        if not sel:
            continue
        for metric in ("f1", "accuracy", "loss"):
            plot(metric, sel, os.path.join(out, f"{gname}_{metric}.png"))
            print("wrote", os.path.join(out, f"{gname}_{metric}.png"))
    with open(os.path.join(out, "LEGEND.txt"), "w") as f:
        for gname, names in groups.items():
            f.write(gname + ": " + ", ".join(n for n in names if n in runs) + "  (colour order: blue, orange, green, red)\n")


if __name__ == "__main__":
    main()
