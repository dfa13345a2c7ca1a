"""Writes the small sample the reference ships as mockData/small-sample/small_sample_reviews_embedded_{training,test}.csv
(stripped from its checkout, .MISSING_LARGE_BLOBS:1-2): same format (header 0..1023,Score), synthetic content — and
mockData/headerless_binary_label.csv (400 rows x 12 real-valued features, label 0 / 1, no header line): the kind of
file the reference's two other mock files are (headerless, binary last column; no code of the reference reads them),
which trains here with --no_header --num_features 12 --num_classes 1.
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
    rng = np.random.default_rng(7)
    X = np.round(rng.standard_normal((400, 12)) * rng.uniform(0.2, 3.0, 12), 3)
    w = rng.standard_normal(12)
    lab = (X @ w + 0.8 * rng.standard_normal(400) > 0).astype(int)
    with open(os.path.join(root, "headerless_binary_label.csv"), "w") as f:
        for r in range(400):
            f.write(",".join(f"{v:g}" for v in X[r]) + f",{lab[r]}\n")
    print("wrote", out, "and", root)


if __name__ == "__main__":
    main()
