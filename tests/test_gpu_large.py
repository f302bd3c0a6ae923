"""Bit-exact parity at BASELINE sizes (tests/golden/large.json).

The digests were produced by the UNMODIFIED reference binaries in the build container
(tests/golden/make_golden.py --large: gsufsort main.c:848-895 writes .bwt/.bwt.qs/.1.lcp, eGap the EOS-0
spelling, bfq_ext decode.cpp:1040-1127 the .fq, bfq_ext.cpp:1113-1128 the counters).  The inputs are
regenerated here from the seeds (tests/large_cases.py), the device outputs are hashed and compared.
"""
import hashlib
import json
import os

import numpy as np
import pytest

import helpers as H
import large_cases as LC
from bfqzip_b200 import dist, lib

pytestmark = pytest.mark.gpu

STAT_KEYS = ["clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"]
PATH = os.path.join(H.GOLDEN, "large.json")
GOLD = json.load(open(PATH)) if os.path.exists(PATH) else {}


def md5(a):
    if hasattr(a, "cpu"):
        a = a.cpu().numpy()
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


_cache = {}


def case(name):
    """(reads, golden record); the last case stays cached (the inputs of a config-4 block are 1.5 GB)."""
    if name not in GOLD:
        pytest.skip(f"{name}: no golden digests in large.json")
    if _cache.get("name") != name:
        _cache.clear()
        reads = LC.make_case(name)
        g = GOLD[name]
        assert md5(reads[0]) == g["in_bases"] and md5(reads[1]) == g["in_quals"], "the seeded generator drifted"
        _cache.update(name=name, reads=reads)
    return _cache["reads"], GOLD[name]


def mode_of(g):
    key = [k for k in g if k.startswith("M") and "_B" in k][0]
    return int(key[1]), int(key[4]), g[key]


def run_dev(name, rebuild, eos):
    """bfq_pipeline_dev with device-resident buffers; returns digests of the five outputs + statistics."""
    import torch
    (bases, quals, off), g = case(name)
    M, B, gm = mode_of(g)
    dev = torch.device("cuda", 0)
    nreads = len(off) - 1
    total = int(off[-1]); n = total + nreads
    c = lib.BfqGpu(M, B, 0)
    try:
        c.set_option("rebuild", rebuild)
        d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        e = lambda k: torch.empty(k + 64, dtype=torch.uint8, device=dev)
        d_bwt, d_qs, d_lcp, d_ob, d_oq = e(n), e(n), e(n), e(total), e(total)
        st = c.pipeline_dev(d(bases), d(quals), d(off.view(np.int64)), nreads, total, lib.default_params(eos=eos),
                            d_bwt, d_qs, d_lcp, d_ob, d_oq)
        out = dict(bwt=md5(d_bwt[:n]), qs=md5(d_qs[:n]), lcp=md5(d_lcp[:n]), out_bases=md5(d_ob[:total]), out_quals=md5(d_oq[:total]))
    finally:
        c.close()
        del c
        torch.cuda.empty_cache()
    return out, st, gm


def check(out, st, g, gm, eos):
    files = g["eos0"] if eos == 0 else g["eos35"]
    assert out["bwt"] == files["bwt"], "eBWT differs from the reference"
    assert out["qs"] == files["qs"], "permuted qualities differ from the reference"
    assert out["lcp"] == files["lcp"], "LCP differs from the reference"
    assert out["out_bases"] == gm["out_bases"], "rebuilt bases differ from the reference .fq"
    assert out["out_quals"] == gm["out_quals"], "rebuilt qualities differ from the reference .fq"
    assert {k: st[k] for k in STAT_KEYS} == gm["stats"]


# ---- config 2 in full: both part-C paths on device buffers, both EOS spellings (eGap's files for 0, gsufsort's for 35)
@pytest.mark.parametrize("rebuild", [1, 2], ids=["unpermute", "lfwalk"])
@pytest.mark.parametrize("eos", [0, 35])
def test_cfg2_full_device(rebuild, eos):
    (_, _, _), g = case("cfg2_full")
    if eos == 0 and "eos0" not in g:
        pytest.skip("no eGap digests")
    out, st, gm = run_dev("cfg2_full", rebuild, eos)
    check(out, st, g, gm, eos)


def test_cfg2_full_host_buffers():
    """bfq_pipeline (host buffers, rows and reads leave slice by slice while later slices run), host_slices 16."""
    (bases, quals, off), g = case("cfg2_full")
    M, B, gm = mode_of(g)
    eos = 0 if "eos0" in g else 35
    c = lib.BfqGpu(M, B, 0)
    try:
        c.set_option("host_slices", 16)
        (bwt, qs, lcp), (ob, oq), st = c.pipeline(bases, quals, off, lib.default_params(eos=eos))
    finally:
        c.close()
    check(dict(bwt=md5(bwt), qs=md5(qs), lcp=md5(lcp), out_bases=md5(ob), out_quals=md5(oq)), st, g, gm, eos)
    text = LC.format_fastq_fast(ob, oq, off)
    assert hashlib.md5(text).hexdigest() == gm["fq"], "the FASTQ text differs from the reference's .fq"


# ---- config 3 (1/8, M=1 B=1), one config-4 block as split_fastq_2 builds it (M=3), both part-C paths
@pytest.mark.parametrize("rebuild", [1, 2], ids=["unpermute", "lfwalk"])
@pytest.mark.parametrize("name", ["cfg3_eighth", "cfg4_block"])
def test_large_device(name, rebuild):
    (_, _, _), g = case(name)
    out, st, gm = run_dev(name, rebuild, 35)
    check(out, st, g, gm, 35)


# ---- config 3 in full (10M x 150 bp, M=1 B=1: the largest single-GPU configuration of BASELINE.json), the default part-C path
def test_cfg3_full_device():
    import psutil
    if psutil.virtual_memory().available < 40 << 30:
        pytest.skip("regenerating 10M x 150 bp on the host takes ~27 GB")
    (_, _, _), g = case("cfg3_full")
    out, st, gm = run_dev("cfg3_full", 1, 35)
    check(out, st, g, gm, 35)
    _cache.clear()          # 3 GB of host arrays


# ---- config 5 (1/8 of the 50M reads, M=0) through the single-collection mode, eight emulated ranks
def test_cfg5_eighth_single_collection_8_ranks():
    reads, g = case("cfg5_eighth")
    M, B, gm = mode_of(g)
    eng = dist.GpuEngine(M, B, 0)
    try:
        bwt, qs, lcp, ob, oq, st = dist.emulate_ranks(eng, reads, lib.default_params(eos=35), 8)
    finally:
        eng.close()
    check(dict(bwt=md5(bwt), qs=md5(qs), lcp=md5(lcp), out_bases=md5(ob), out_quals=md5(oq)), st, g, gm, 35)
