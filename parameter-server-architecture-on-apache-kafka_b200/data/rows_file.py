"""PSBROWS1: a data set converted once into the palette wire format of ingestion.

The reference streams CSV text (``CsvProducer.java:36-87``); tokenising text is what bounds a host thread (about
0.5 M rows/s on the GPU box's CPU, ``profiles/README.md``) while one B200 trains on 0.6 M fresh rows/s and more.
A rows file holds exactly what the producer thread would have produced from the CSV — presence bitmap, per-row
palette + 4-bit codes (``psb_pack_palette_dense``), label — so streaming it is two copies per row and nothing to
parse.  Every entry point that takes a CSV path takes a rows file as well (the producer thread looks at the magic).

Layout (little-endian; every section starts on a 64-byte boundary):
    header   64 B   magic "PSBROWS1", u32 num_features, u32 version = 1, u64 rows, u64 payload bytes
    offsets  u64[rows + 1]    row i's payload is payload[offsets[i] : offsets[i + 1]]
    labels   i32[rows]
    bitmap   u32[rows][num_features / 32]
    payload  bytes            per row: u16 npal; npal > 0: u16 palette[npal] + 4-bit codes (padded to even);
                              npal == 0: the fp16 values of the row's non-zeros
"""
from __future__ import annotations

import struct

import numpy as np

from .csr import CSRBatch

MAGIC = b"PSBROWS1"


def _up64(x: int) -> int:
    return (x + 63) & ~63


def is_rows_file(path: str) -> bool:
    try:
        with open(path, "rb") as f:
            return f.read(8) == MAGIC
    except OSError:
        return False


def write_rows_file(path: str, batch: CSRBatch, chunk: int = 8192) -> int:
    """Packs `batch` (any num_features that is a multiple of 32) and writes it; returns the file size."""
    from ..ops import native
    F, n = batch.num_features, len(batch)
    if F % 32 != 0:
        raise ValueError("a rows file needs num_features % 32 == 0")
    words = F // 32
    offsets = np.zeros(n + 1, dtype=np.uint64)
    bitmaps, payloads = [], []
    pk = native.PackedRows(min(chunk, max(n, 1)), F, pin=False, palette=True)
    total = 0
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        part = CSRBatch(batch.indptr[a:b + 1] - batch.indptr[a], batch.indices[batch.indptr[a]:batch.indptr[b]],
                        batch.values[batch.indptr[a]:batch.indptr[b]], batch.labels[a:b], F)
        pk.pack_csr(part)
        offsets[a:b + 1] = pk.row_off[: b - a + 1].numpy().astype(np.uint64) + np.uint64(total)
        bitmaps.append(pk.bitmap[: b - a].numpy().copy())
        payloads.append(pk.payload[: pk.payload_bytes].numpy().copy())
        total += pk.payload_bytes
    with open(path, "wb") as f:
        f.write(MAGIC + struct.pack("<IIQQ", F, 1, n, total) + b"\0" * 32)
        at = 64
        for arr, size in ((offsets, (n + 1) * 8), (batch.labels.astype(np.int32), n * 4),
                          (np.concatenate(bitmaps) if bitmaps else np.zeros((0, words), np.int32), n * words * 4)):
            f.write(np.ascontiguousarray(arr).tobytes())
            f.write(b"\0" * (_up64(at + size) - (at + size)))
            at = _up64(at + size)
        for p in payloads:
            f.write(p.tobytes())
        return at + total


def read_rows_file(path: str) -> CSRBatch:
    """The whole file as a CSR batch (test sets, in-memory producers, tests).  Values are the stored fp16 values."""
    raw = np.fromfile(path, dtype=np.uint8)
    if raw[:8].tobytes() != MAGIC:
        raise ValueError(f"{path} is not a PSBROWS1 file")
    F, version, n, total = struct.unpack("<IIQQ", raw[8:32].tobytes())
    if version != 1 or F % 32 != 0:
        raise ValueError(f"{path}: unsupported rows file (version {version}, {F} features)")
    words, at = F // 32, 64
    off = raw[at:at + (n + 1) * 8].view(np.uint64).astype(np.int64)
    at = _up64(at + (n + 1) * 8)
    labels = raw[at:at + n * 4].view(np.int32).copy()
    at = _up64(at + n * 4)
    bitmap = raw[at:at + n * words * 4].view(np.uint32).reshape(n, words)
    at = _up64(at + n * words * 4)
    pay = raw[at:at + total]
    bits = np.unpackbits(bitmap.view(np.uint8), axis=1, bitorder="little").astype(bool)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(bits.sum(axis=1), out=indptr[1:])
    indices = np.nonzero(bits)[1].astype(np.int32)
    values = np.zeros(int(indptr[-1]), dtype=np.float16)
    for r in range(n):
        p = pay[off[r]: off[r + 1]]
        nnz = int(indptr[r + 1] - indptr[r])
        npal = int(p[:2].view(np.uint16)[0])
        if npal == 0:
            vals = p[2: 2 + 2 * nnz].view(np.float16)
        else:
            pal = p[2: 2 + 2 * npal].view(np.float16)
            c = p[2 + 2 * npal: 2 + 2 * npal + (nnz + 1) // 2]
            vals = pal[np.stack([c & 15, c >> 4], axis=1).reshape(-1)[:nnz]]
        values[indptr[r]: indptr[r + 1]] = vals
    return CSRBatch(indptr, indices, values.astype(np.float32), labels, F)


__all__ = ["MAGIC", "is_rows_file", "write_rows_file", "read_rows_file"]
