"""BASELINE-size parity cases (tests/golden/large.json): seeded inputs shared by the golden generator
(tests/golden/make_golden.py --large, run with the reference binaries in the build container) and the GPU tests.

Every case is a fixed-length collection, so FASTQ text is produced and parsed with numpy reshapes.
"""
from __future__ import annotations

import numpy as np

from bfqzip_b200 import parallel, synth

# BASELINE.json configs 2-5 (SURVEY 8(d) generator parameters).  cfg3/cfg5 are 1/8 of the named size at the same
# 40x coverage; cfg4 is ONE of the eight blocks, cut the way split_fastq_2 (BFQzip_parallel.py:325-360) builds it:
# mates-1 of the block followed by mates-2 of the block, 5x coverage of the 150 Mbp genome per block.
CASES = {
    "cfg2_full": dict(kind="single", nreads=1_000_000, L=100, seed=1001, coverage=40.0, M=2, B=0, egap=True),
    "cfg3_eighth": dict(kind="single", nreads=1_250_000, L=150, seed=1002, coverage=40.0, M=1, B=1),
    "cfg3_full": dict(kind="single", nreads=10_000_000, L=150, seed=1002, coverage=40.0, M=1, B=1),
    "cfg4_block": dict(kind="paired", npairs=2_500_000, L=150, seed=1003, coverage=5.0, M=3, B=0),
    "cfg5_eighth": dict(kind="single", nreads=6_250_000, L=150, seed=1004, coverage=40.0, M=0, B=0),
}


def make_case(name):
    c = CASES[name]
    if c["kind"] == "single":
        return synth.make_reads(c["nreads"], c["L"], c["seed"], coverage=c["coverage"])
    r1, r2 = synth.make_paired(c["npairs"], c["L"], c["seed"], coverage=c["coverage"])
    blocks = parallel.make_blocks(r1, 1, r2)
    assert len(blocks) == 1 and blocks[0][1] == c["npairs"]
    return blocks[0][0]


def format_fastq_fast(bases, quals, off) -> bytes:
    """'@\\nBASES\\n+\\nQUALS\\n' records (decode.cpp:1080-1127) for a fixed-length collection."""
    R = len(off) - 1
    L = int(off[1])
    assert int(off[-1]) == R * L
    rec = np.empty((R, 2 * L + 6), np.uint8)
    rec[:, 0] = ord("@"); rec[:, 1] = 10
    rec[:, 2:2 + L] = bases.reshape(R, L)
    rec[:, 2 + L] = 10; rec[:, 3 + L] = ord("+"); rec[:, 4 + L] = 10
    rec[:, 5 + L:5 + 2 * L] = quals.reshape(R, L)
    rec[:, 5 + 2 * L] = 10
    return rec.tobytes()


def parse_fastq_fast(text: np.ndarray, off):
    R = len(off) - 1
    L = int(off[1])
    assert text.size == R * (2 * L + 6), (text.size, R, L)
    rec = text.reshape(R, 2 * L + 6)
    assert (rec[:, 0] == ord("@")).all() and (rec[:, 1] == 10).all() and (rec[:, 3 + L] == ord("+")).all()
    return np.ascontiguousarray(rec[:, 2:2 + L]).reshape(-1), np.ascontiguousarray(rec[:, 5 + L:5 + 2 * L]).reshape(-1)
