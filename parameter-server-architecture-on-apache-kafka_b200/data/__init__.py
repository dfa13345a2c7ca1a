from .csr import CSRBatch, pack_stage_rows, unpack_stage_rows, stage_stride_words
from .csv_io import read_csv, write_csv, iter_rows, parse_line, header
from .synthetic import ReviewGenerator, random_csr, TEST_SUPPORTS

__all__ = ["CSRBatch", "pack_stage_rows", "unpack_stage_rows", "stage_stride_words", "read_csv",
           "write_csv", "iter_rows", "parse_line", "header", "ReviewGenerator", "random_csr",
           "TEST_SUPPORTS"]
