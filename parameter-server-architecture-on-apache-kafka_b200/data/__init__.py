from .csr import CSRBatch, pack_stage_rows, unpack_stage_rows, stage_stride_words
from .csv_io import read_csv, write_csv, iter_rows, parse_line, header
from .rows_file import is_rows_file, read_rows_file, write_rows_file
from .synthetic import ReviewGenerator, random_csr, TEST_SUPPORTS

__all__ = ["CSRBatch", "pack_stage_rows", "unpack_stage_rows", "stage_stride_words", "read_csv",
           "write_csv", "iter_rows", "parse_line", "header", "ReviewGenerator", "random_csr",
           "TEST_SUPPORTS", "is_rows_file", "read_rows_file", "write_rows_file"]
