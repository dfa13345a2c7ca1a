"""Writes the small sample the reference ships as mockData/small-sample/small_sample_reviews_embedded_{training,test}.csv
(stripped from its checkout, .MISSING_LARGE_BLOBS:1-2): same format (header 0..1023,Score), synthetic content.
Usage: python tools/make_mock_data.py [train_rows test_rows out_dir]"""
import os
import sys

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
    print("wrote", out)


if __name__ == "__main__":
    main()
