"""CSV -> PSBROWS1 rows file: the data set converted once into the palette wire format of ingestion, so that streaming
it costs two copies per row instead of tokenising ~2.3 KB of text (psb200/data/rows_file.py).  Every -training / -test
option takes the result in place of the CSV.

    python tools/csv_to_rows.py train.csv train.psbrows [--no_header]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from psb200.data import read_csv, write_rows_file
from psb200.ops import native


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("out")
    ap.add_argument("--no_header", action="store_true")
    a = ap.parse_args()
    t0 = time.time()
    batch = native.csv_to_csr(a.csv, has_header=not a.no_header) if native.available() else \
        read_csv(a.csv, has_header=not a.no_header)
    if batch.num_features % 32:
        sys.exit(f"{a.csv}: {batch.num_features} features; a rows file needs a multiple of 32")
    size = write_rows_file(a.out, batch)
    print(f"{len(batch)} rows x {batch.num_features} features: {os.path.getsize(a.csv) / 1e6:.1f} MB of text -> "
          f"{size / 1e6:.1f} MB ({size / max(1, len(batch)):.0f} B per row) in {time.time() - t0:.1f} s")


if __name__ == "__main__":
    main()
