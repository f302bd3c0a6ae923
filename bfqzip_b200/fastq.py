"""FASTQ <-> flat read arrays (host side of the boundary).

Mirrors what the reference's loaders keep of a FASTQ file: line 2 (bases) and line 4
(qualities) of every 4-line record, in file order (external/gsufsort/lib/file.c:182-275),
and what its writer emits: ``header\\nBASES\\n+\\nQUALS\\n`` with header ``@`` unless a
header file is given (src_ext_mem/decode.cpp:1040-1127).
"""
from __future__ import annotations

import numpy as np


def parse_fastq_bytes(data: bytes):
    """Returns (bases u8, quals u8, off u64[nreads+1], headers list[bytes])."""
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    nrec = len(lines) // 4
    headers = lines[0:4 * nrec:4]
    seqs = [s[:-1] if s.endswith(b"\r") else s for s in lines[1:4 * nrec:4]]       # CR before LF is not part of the read
    quals = [s[:-1] if s.endswith(b"\r") else s for s in lines[3:4 * nrec:4]]
    lens = np.fromiter((len(s) for s in seqs), dtype=np.int64, count=nrec)
    off = np.zeros(nrec + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:].view(np.int64))
    bases = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy()
    q = np.frombuffer(b"".join(quals), dtype=np.uint8).copy()
    if q.size != bases.size:
        raise ValueError("FASTQ: bases and qualities differ in length")
    return bases, q, off, headers


def read_fastq(path: str):
    with open(path, "rb") as f:
        return parse_fastq_bytes(f.read())


def format_fastq(bases: np.ndarray, quals: np.ndarray, off: np.ndarray, headers=None) -> bytes:
    """4-line records exactly as decode.cpp:1080-1127 writes them."""
    out = []
    b = bases.tobytes()
    q = quals.tobytes()
    o = off.astype(np.int64)
    for g in range(len(o) - 1):
        h = headers[g] if headers is not None else b"@"
        out.append(h + b"\n" + b[o[g]:o[g + 1]] + b"\n+\n" + q[o[g]:o[g + 1]] + b"\n")
    return b"".join(out)


def write_fastq(path: str, bases, quals, off, headers=None) -> None:
    with open(path, "wb") as f:
        f.write(format_fastq(bases, quals, off, headers))


def format_fastq_fixed(bases: np.ndarray, quals: np.ndarray, nreads: int, read_len: int) -> bytes:
    """format_fastq for a fixed-length collection with '@' headers, as one numpy reshape (millions of reads per second)."""
    rec = np.empty((nreads, 2 * read_len + 6), np.uint8)
    rec[:, 0] = ord("@"); rec[:, 1] = 10
    rec[:, 2:2 + read_len] = np.asarray(bases, np.uint8).reshape(nreads, read_len)
    rec[:, 2 + read_len] = 10; rec[:, 3 + read_len] = ord("+"); rec[:, 4 + read_len] = 10
    rec[:, 5 + read_len:5 + 2 * read_len] = np.asarray(quals, np.uint8).reshape(nreads, read_len)
    rec[:, 5 + 2 * read_len] = 10
    return rec.tobytes()
