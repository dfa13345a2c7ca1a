"""Writes the small sample the reference ships as mockData/small-sample/small_sample_reviews_embedded_{training,test}.csv
(stripped from its checkout, .MISSING_LARGE_BLOBS:1-2): same format (header 0..1023,Score), synthetic content — and
synthetic stand-ins of the shape of its two other mock files (no code of the reference reads them; here both train
with --no_header): mockData/lr_dataset_stripped.csv (570 rows: year, month, day-of-cycle, two small integer
covariates, binary label; no header) and mockData/sample_input_data.csv (49 rows x 100 columns of 0 / 1, the last
one the label; no header).
Usage: python tools/make_mock_data.py [train_rows test_rows out_dir]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from psb200.data import ReviewGenerator, write_csv


def main():
    n_train = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    n_test = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                               "mockData", "small-sample")
    os.makedirs(out, exist_ok=True)
    gen = ReviewGenerator(mean_len=420.0)
    write_csv(os.path.join(out, "small_sample_reviews_embedded_training.csv"), gen.train_split(n_train))
    write_csv(os.path.join(out, "small_sample_reviews_embedded_test.csv"), gen.test_split(n_test))
    root = os.path.dirname(out.rstrip(os.sep))
    rng = np.random.default_rng(1983)
    with open(os.path.join(root, "lr_dataset_stripped.csv"), "w") as f:      # 38 (year, month) cycles of 15 samples
        for i in range(570):
            cycle, k = divmod(i, 15)
            year, month = 1983 + (cycle + 2) // 12, (cycle + 2) % 12 + 1
            day = 0 if k == 0 else 3 * k - 1
            group = cycle % 4 + 1
            flag = int(rng.random() < 0.1)
            p = 1.0 / (1.0 + np.exp(-(0.9 * (group - 2.5) - 0.04 * (day - 20) + 1.5 * flag)))
            f.write(f"{year},{month},{day},{flag},{group},{int(rng.random() < p)}\n")
    bits = (rng.random((49, 99)) < 0.5).astype(int)
    w = rng.standard_normal(99)
    lab = (bits @ w + 0.5 * rng.standard_normal(49) > np.median(bits @ w)).astype(int)
    with open(os.path.join(root, "sample_input_data.csv"), "w") as f:
        for r in range(49):
            f.write(",".join(map(str, bits[r])) + f",{lab[r]}\n")
    print("wrote", out, "and", root)


if __name__ == "__main__":
    main()
