"""Pins the oracle (oracle/oracle.c) against the reference's own outputs.

* golden: tests/golden/*.npz were produced by the unmodified reference binaries
  (tests/golden/make_golden.py); this runs everywhere, including the GPU box.
* live: when oracle/_ref/ exists (built from /root/reference by oracle/Makefile), fresh random
  inputs are pushed through both the reference binaries and the oracle.
"""
import glob
import os
import tempfile

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import fastq, synth

MODES = [(M, B) for M in range(4) for B in range(2)]
STAT_KEYS = ["clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"]
GOLD = sorted(glob.glob(os.path.join(H.GOLDEN, "*.npz")))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_oracle_part_a_golden(path):
    g = np.load(path)
    bwt, qs, lcp = H.oracle_build_ebwt(g["bases"], g["quals"], g["off"], 35)
    assert np.array_equal(bwt, g["bwt"])
    assert np.array_equal(qs, g["qs"])
    assert np.array_equal(lcp, g["lcp"])


@pytest.mark.parametrize("mode", MODES, ids=[f"M{m}B{b}" for m, b in MODES])
@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p)[:-4] for p in GOLD])
def test_oracle_part_bc_golden(path, mode):
    M, B = mode
    g = np.load(path)
    p = H.make_params(M=M, B=B)
    b, q, off, st, _, _, err = H.oracle_smooth_rebuild(g["bwt"], g["qs"], g["lcp"], p)
    assert err == 0
    assert np.array_equal(off, g["off"])
    assert np.array_equal(b, g[f"out_bases_M{M}_B{B}"])
    assert np.array_equal(q, g[f"out_quals_M{M}_B{B}"])
    assert [st[k] for k in STAT_KEYS] == [int(x) for x in g[f"stats_M{M}_B{B}"]]


def test_known_answer_example():
    """example/reads.fastq, M=2 B=0 -> 387 clusters, 4198 qs, 10 bases."""
    import hashlib
    g = np.load(os.path.join(H.GOLDEN, "example.npz"))
    p = H.make_params(M=2, B=0)
    b, q, off, st, _, _, _ = H.oracle_smooth_rebuild(g["bwt"], g["qs"], g["lcp"], p)
    headers = g["headers"].tobytes().split(b"\n")
    fq = fastq.format_fastq(b, q, off, headers)
    assert hashlib.md5(fq).hexdigest() == "9178301c8ef6c9864d5ebfccf47313d4"
    assert (st["clusters"], st["qs_smoothed"], st["modified"]) == (387, 4198, 10)
    sym, cnt = np.unique(g["bwt"], return_counts=True)
    assert dict(zip(sym.tolist(), cnt.tolist())) == {35: 100, 65: 2876, 67: 2077, 71: 2294, 78: 2, 84: 2851}


def test_binning_table():
    lib = H.oracle_lib()
    exp = {0: 0, 1: 1, 2: 6, 9: 6, 10: 15, 19: 15, 20: 22, 24: 22, 25: 27, 29: 27, 30: 33, 34: 33,
           35: 37, 39: 37, 40: 40, 41: 40, 60: 40}
    for q, e in exp.items():
        assert lib.oracle_binning(q + 33) == e + 33


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built (no /root/reference here)")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_oracle_vs_live_reference(seed):
    rng = np.random.default_rng(seed)
    nreads = int(rng.integers(200, 3000)); L = int(rng.integers(30, 120))
    bases, quals, off = synth.make_reads(nreads, L, 1000 + seed, coverage=float(rng.integers(10, 80)),
                                         var_len=bool(seed & 1), err=0.02)
    wd = tempfile.mkdtemp()
    fq = os.path.join(wd, "in.fastq")
    fastq.write_fastq(fq, bases, quals, off)
    rb, rq, rl = H.ref_gsufsort(fq, wd)
    ob, oq, ol = H.oracle_build_ebwt(bases, quals, off, 35)
    assert np.array_equal(rb, ob) and np.array_equal(rq, oq) and np.array_equal(rl, ol)
    for M, B in [(seed % 4, 0), ((seed + 1) % 4, 1)]:
        p = H.make_params(M=M, B=B)
        b, q, o, st, _, _, _ = H.oracle_smooth_rebuild(ob, oq, ol, p)
        ref, rst = H.ref_bfq_ext(rb, rq, rl, M=M, B=B)
        assert fastq.format_fastq(b, q, o) == ref
        assert all(st[k] == rst[k] for k in rst)
