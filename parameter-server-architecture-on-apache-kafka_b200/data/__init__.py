from .csr import CSRBatch, pack_rows, unpack_rows, slot_words, NOWN, HDR_WORDS
from .csv_io import read_csv, write_csv, iter_rows, parse_line, header
from .synthetic import ReviewGenerator, random_csr, TEST_SUPPORTS

__all__ = ["CSRBatch", "pack_rows", "unpack_rows", "slot_words", "NOWN", "HDR_WORDS", "read_csv",
           "write_csv", "iter_rows", "parse_line", "header", "ReviewGenerator", "random_csr",
           "TEST_SUPPORTS"]
