"""Test-only helpers: ctypes binding of oracle/liboracle.so and a runner for oracle/_ref.

Nothing here is imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")


class OracleParams(C.Structure):
    _fields_ = [("K", C.c_int32), ("m", C.c_int32), ("qthr", C.c_int32), ("max_len", C.c_int32),
                ("M", C.c_int32), ("B", C.c_int32), ("fthr", C.c_float),
                ("eos", C.c_uint8), ("default_q", C.c_uint8)]


class OracleStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("clusters", "discarded", "ambiguous", "processed",
                                          "bases_inside", "qs_smoothed", "modified", "nreads")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def make_params(M=2, B=0, K=16, m=5, qthr=20, max_len=250, fthr=40.0, eos=35, default_q=ord(">")):
    return OracleParams(K, m, qthr, max_len, M, B, fthr, eos, default_q)


_lib = None


def oracle_lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "liboracle.so")
        src = os.path.join(ORACLE_DIR, "oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "port"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(so)
        _lib.oracle_build_ebwt.restype = C.c_uint64
        _lib.oracle_rebuild.restype = C.c_uint64
        _lib.oracle_newqs.restype = C.c_uint8
        _lib.oracle_binning.restype = C.c_uint8
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_build_ebwt(bases, quals, off, eos=35):
    lib = oracle_lib()
    nreads = len(off) - 1
    n = int(off[-1]) + nreads
    bwt = np.empty(n, np.uint8); qs = np.empty(n, np.uint8); lcp = np.empty(n, np.uint8)
    bases = np.ascontiguousarray(bases, np.uint8); quals = np.ascontiguousarray(quals, np.uint8)
    off = np.ascontiguousarray(off, np.uint64)
    got = lib.oracle_build_ebwt(_p(bases), _p(quals), _p(off), C.c_uint64(nreads), C.c_uint8(eos),
                                _p(bwt), _p(qs), _p(lcp))
    assert got == n
    return bwt, qs, lcp


def oracle_smooth(bwt, qs, lcp, params):
    lib = oracle_lib()
    n = len(bwt)
    sq = np.empty(n, np.uint8); ob = np.empty(n, np.uint8)
    st = OracleStats()
    err = lib.oracle_smooth(_p(bwt), _p(qs), _p(lcp), C.c_uint64(n), C.byref(params),
                            _p(sq), _p(ob), C.byref(st))
    return sq, ob, st.as_dict(), err


def oracle_rebuild(bwt, obwt, sq, params, nreads):
    lib = oracle_lib()
    n = len(bwt)
    ob = np.empty(max(1, n - nreads), np.uint8); oq = np.empty(max(1, n - nreads), np.uint8)
    off = np.empty(nreads + 1, np.uint64)
    w = lib.oracle_rebuild(_p(bwt), _p(obwt), _p(sq), C.c_uint64(n), C.byref(params),
                           C.c_uint64(nreads), _p(ob), _p(oq), _p(off))
    return ob[:w], oq[:w], off


def oracle_smooth_rebuild(bwt, qs, lcp, params):
    sq, ob, st, err = oracle_smooth(bwt, qs, lcp, params)
    b, q, off = oracle_rebuild(bwt, ob, sq, params, st["nreads"])
    return b, q, off, st, sq, ob, err


# ------------------------------------------------------------------ reference binaries
def have_ref():
    return os.path.exists(os.path.join(REF_DIR, "gsufsort")) and \
        os.path.exists(os.path.join(REF_DIR, "bfq_ext_M2_B0"))


def ref_gsufsort(fastq_path, workdir):
    """gsufsort IN --bwt --qs --lcp 1 -o OUT (BFQzip.py:184 spelling + --lcp).  EOS byte 35."""
    out = os.path.join(workdir, "out")
    subprocess.check_call([os.path.join(REF_DIR, "gsufsort"), fastq_path, "--bwt", "--qs",
                           "--lcp", "1", "-o", out], stdout=subprocess.DEVNULL,
                          stderr=subprocess.DEVNULL)
    rd = lambda e: np.fromfile(out + e, dtype=np.uint8)
    return rd(".bwt"), rd(".bwt.qs"), rd(".1.lcp")


def parse_ref_stats(log: str):
    import re
    st = {}
    st["clusters"] = int(re.search(r"(\d+) is the total number of clusters", log).group(1))
    st["discarded"] = int(re.search(r"Discarded: (\d+)\(", log).group(1))
    st["ambiguous"] = int(re.search(r"Ambiguous discarded: (\d+)\(", log).group(1))
    st["processed"] = int(re.search(r"Processed: (\d+)\(", log).group(1))
    st["bases_inside"] = int(re.search(r"(\d+)/\d+ \([^)]*\) bases fall inside", log).group(1))
    st["qs_smoothed"] = int(re.search(r"(\d+)/\d+ qualities have been modified", log).group(1))
    st["modified"] = int(re.search(r"(\d+)/\d+ bases have been modified", log).group(1))
    return st


def ref_bfq_ext(bwt, qs, lcp, M=2, B=0, eos=35, K=None, m=5, v=None, headers=None, threads=None):
    """Runs oracle/_ref/bfq_ext_M?_B? in a private cwd (it writes ./tmp).  Returns (fq bytes, stats)."""
    wd = tempfile.mkdtemp(prefix="bfqref_")
    try:
        for name, arr in (("x.bwt", bwt), ("x.bwt.qs", qs), ("x.1.lcp", lcp)):
            np.asarray(arr, np.uint8).tofile(os.path.join(wd, name))
        cmd = [os.path.join(REF_DIR, f"bfq_ext_M{M}_B{B}"), "-e", "x.bwt", "-q", "x.bwt.qs",
               "-a", "x.1.lcp", "-o", "x", "-l", "250", "-s", str(eos), "-m", str(m)]
        if K is not None:
            cmd += ["-k", str(K)]
        if v is not None:
            cmd += ["-v", str(v)]
        if threads is not None:
            cmd += ["-T", str(threads)]
        if headers is not None:
            with open(os.path.join(wd, "x.h"), "wb") as f:
                f.write(b"".join(h + b"\n" for h in headers))
            cmd += ["-H", "x.h"]
        r = subprocess.run(cmd, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
        if r.returncode != 0:
            raise RuntimeError(f"reference bfq_ext failed rc={r.returncode}")
        with open(os.path.join(wd, "x.fq"), "rb") as f:
            fq = f.read()
        return fq, parse_ref_stats(r.stdout.decode("latin1"))
    finally:
        shutil.rmtree(wd, ignore_errors=True)


# ------------------------------------------------------------------ quality-stream pre-transform
def rle_encode_numpy(stream: np.ndarray):
    """CPU restatement of bfq_qs_rle_encode, numpy: maximal runs of equal bytes -> (values, lengths).  Test infrastructure:
    there is no counterpart in the reference (its step 5 takes the stream as it is, BFQzip_ext.py:248-272), so the checker
    is this restatement plus the round trip."""
    stream = np.ascontiguousarray(stream, np.uint8)
    if stream.size == 0:
        return np.empty(0, np.uint8), np.empty(0, np.uint8)
    heads = np.flatnonzero(np.concatenate([[True], stream[1:] != stream[:-1]]))
    lens = np.diff(np.concatenate([heads, [stream.size]]))
    assert lens.max() <= 255
    return stream[heads], lens.astype(np.uint8)
