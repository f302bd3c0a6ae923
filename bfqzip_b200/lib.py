"""ctypes binding of the C ABI in include/bfq_gpu.h (one shared object per compile-time mode).

This is plumbing for tests and bench.py; the product is the CUDA library.  There is no CPU
fallback: if the shared object is missing or no CUDA device is present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBDIR = os.path.join(_HERE, "lib")


class BfqError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"bfq error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    """bfq_params: the run-time flags of bfq_ext (bfq_ext.cpp:105-129)."""
    _fields_ = [("K", C.c_int32), ("m", C.c_int32), ("qthr", C.c_int32), ("max_len", C.c_int32),
                ("fthr", C.c_float), ("eos_byte", C.c_uint8), ("default_q", C.c_uint8)]


class Stats(C.Structure):
    """bfq_stats: the statistics block bfq_ext prints (bfq_ext.cpp:1113-1128)."""
    _fields_ = [(k, C.c_uint64) for k in ("clusters", "discarded", "ambiguous", "processed",
                                          "bases_inside", "qs_smoothed", "modified", "nreads", "n")]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


def lib_path(M: int, B: int) -> str:
    return os.path.join(LIBDIR, f"libbfqgpu_M{M}_B{B}.so")


PROF_CLASSES = ("radix_pass", "radix_hist", "pack", "keys", "group", "defer", "hist", "rank",
                "cluster", "cluster_long", "rule11", "walk", "assemble", "unpermute")

_libs = {}


def load(M: int = 2, B: int = 0) -> C.CDLL:
    key = (M, B)
    if key in _libs:
        return _libs[key]
    path = lib_path(M, B)
    if not os.path.exists(path):
        raise FileNotFoundError(
            f"{path} is missing: build it with `make -C bfqzip_b200/csrc` (or __graft_entry__.build()). "
            "There is no CPU fallback.")
    lib = C.CDLL(path)
    lib.bfq_last_error.restype = C.c_char_p
    lib.bfq_launch_count.restype = C.c_uint64
    lib.bfq_launch_count.argtypes = [C.c_void_p]
    assert lib.bfq_mode_M() == M and lib.bfq_mode_B() == B
    _libs[key] = lib
    return lib


def _ptr(a):
    if a is None:
        return C.c_void_p(0)
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):          # torch tensor (host pinned or device)
        return C.c_void_p(a.data_ptr())
    if isinstance(a, int):
        return C.c_void_p(a)
    raise TypeError(type(a))


def default_params(eos=35, m=5, K=16, qthr=20, max_len=250, fthr=40.0, default_q=ord(">")) -> Params:
    """Driver-faithful defaults: BFQzip_ext.py:207-215 passes -l 250 -m 5; K, t, f, v default."""
    return Params(K, m, qthr, max_len, fthr, eos, default_q)


class BfqGpu:
    """One context = one GPU + one stream (bfq_ctx)."""

    def __init__(self, M: int = 2, B: int = 0, device: int = 0):
        self.M, self.B = M, B
        self.lib = load(M, B)
        h = C.c_void_p()
        rc = self.lib.bfq_ctx_create(C.byref(h), C.c_int(device))
        if rc != 0:
            raise BfqError(rc, self.lib.bfq_last_error().decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.bfq_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def _chk(self, rc):
        if rc != 0:
            raise BfqError(rc, self.lib.bfq_last_error().decode())

    def set_option(self, key: str, value: int):
        self._chk(self.lib.bfq_ctx_set_option(self.h, key.encode(), C.c_int64(value)))

    def set_stream(self, cuda_stream_ptr: int | None):
        self._chk(self.lib.bfq_ctx_set_stream(self.h, C.c_void_p(cuda_stream_ptr or 0)))

    def launch_count(self) -> int:
        return int(self.lib.bfq_launch_count(self.h))

    def prof_reset(self):
        self.lib.bfq_prof_reset(self.h)

    def prof_get(self, klass: str):
        ms = C.c_double(); n = C.c_uint64(); b = C.c_uint64()
        self._chk(self.lib.bfq_prof_get(self.h, klass.encode(), C.byref(ms), C.byref(n), C.byref(b)))
        return ms.value, int(n.value), int(b.value)

    # ---- part A
    def build_ebwt(self, bases, quals, off, eos=35):
        off = np.ascontiguousarray(off, np.uint64)
        nreads = len(off) - 1
        n = int(off[-1]) + nreads
        bwt = np.empty(n, np.uint8); qs = np.empty(n, np.uint8); lcp = np.empty(n, np.uint8)
        self._chk(self.lib.bfq_build_ebwt(self.h, _ptr(bases), _ptr(quals), _ptr(off), C.c_uint64(nreads),
                                          C.c_uint8(eos), _ptr(bwt), _ptr(qs), _ptr(lcp)))
        return bwt, qs, lcp

    # ---- parts B+C
    def smooth_rebuild(self, bwt, qs, lcp, params: Params):
        n = len(bwt)
        nreads = int(np.count_nonzero(np.asarray(bwt) == params.eos_byte))
        ob = np.empty(max(n, 1), np.uint8); oq = np.empty(max(n, 1), np.uint8)
        off = np.empty(nreads + 1, np.uint64)
        st = Stats()
        self._chk(self.lib.bfq_smooth_rebuild(self.h, _ptr(bwt), _ptr(qs), _ptr(lcp), C.c_uint64(n),
                                              C.byref(params), _ptr(ob), _ptr(oq), _ptr(off),
                                              C.c_uint64(nreads + 1), C.byref(st)))
        total = int(off[-1])
        return ob[:total], oq[:total], off, st.as_dict()

    def export_rows(self, n):
        q = np.empty(n, np.uint8); b = np.empty(n, np.uint8)
        self._chk(self.lib.bfq_export_rows(self.h, _ptr(q), _ptr(b)))
        return q, b

    # ---- fused, host buffers (numpy arrays or pinned torch tensors)
    def pipeline(self, bases, quals, off, params: Params, want_files=True, out=None):
        off_np = off if isinstance(off, np.ndarray) else off.numpy()
        nreads = len(off_np) - 1
        total = int(off_np[-1]); n = total + nreads
        if out is None:
            bwt = np.empty(n, np.uint8) if want_files else None
            qs = np.empty(n, np.uint8) if want_files else None
            lcp = np.empty(n, np.uint8) if want_files else None
            ob = np.empty(max(total, 1), np.uint8); oq = np.empty(max(total, 1), np.uint8)
        else:
            bwt, qs, lcp, ob, oq = out
        st = Stats()
        self._chk(self.lib.bfq_pipeline(self.h, _ptr(bases), _ptr(quals), _ptr(off), C.c_uint64(nreads),
                                        C.byref(params), _ptr(bwt), _ptr(qs), _ptr(lcp), _ptr(ob), _ptr(oq),
                                        C.byref(st)))
        return (bwt, qs, lcp), (ob[:total], oq[:total]), st.as_dict()

    # ---- fused, device buffers (torch CUDA tensors or raw device pointers)
    def pipeline_dev(self, d_bases, d_quals, d_off, nreads, total_bases, params: Params,
                     d_bwt, d_qs, d_lcp, d_out_bases, d_out_quals):
        st = Stats()
        self._chk(self.lib.bfq_pipeline_dev(self.h, _ptr(d_bases), _ptr(d_quals), _ptr(d_off),
                                            C.c_uint64(nreads), C.c_uint64(total_bases), C.byref(params),
                                            _ptr(d_bwt), _ptr(d_qs), _ptr(d_lcp), _ptr(d_out_bases),
                                            _ptr(d_out_quals), C.byref(st)))
        return st.as_dict()

    # ---- FASTQ text on either side (parsed / written on the device)
    def build_ebwt_fastq(self, text: bytes, eos=35):
        buf = np.frombuffer(text, np.uint8)
        cap = max(1, len(buf) // 2)
        bwt = np.empty(cap, np.uint8); qs = np.empty(cap, np.uint8); lcp = np.empty(cap, np.uint8)
        nreads = C.c_uint64(0); n = C.c_uint64(0)
        self._chk(self.lib.bfq_build_ebwt_fastq(self.h, _ptr(buf), C.c_uint64(len(buf)), C.c_uint8(eos), _ptr(bwt), _ptr(qs),
                                                _ptr(lcp), C.c_uint64(cap), C.byref(nreads), C.byref(n)))
        k = int(n.value)
        return bwt[:k], qs[:k], lcp[:k], int(nreads.value)

    def smooth_rebuild_fastq(self, bwt, qs, lcp, params: Params, components=False):
        n = len(bwt)
        dna = np.empty(n + 1, np.uint8) if components else None
        qsl = np.empty(n + 1, np.uint8) if components else None
        cap = 6 * n + 64                       # 2 T + 6 R bytes for T bases in R reads, n = T + R
        out = np.empty(cap, np.uint8)
        size = C.c_uint64(0); st = Stats()
        self._chk(self.lib.bfq_smooth_rebuild_fastq(self.h, _ptr(bwt), _ptr(qs), _ptr(lcp), C.c_uint64(n), C.byref(params),
                                                    _ptr(out), C.c_uint64(cap), C.byref(size), _ptr(dna), _ptr(qsl), C.byref(st)))
        text = out[:int(size.value)].tobytes()
        if components:
            r = text.count(b"\n") // 4
            k = (len(text) - 6 * r) // 2 + r                      # bases + one newline per read
            return text, st.as_dict(), dna[:k].tobytes(), qsl[:k].tobytes()
        return text, st.as_dict()

    def pipeline_fastq(self, text: bytes, params: Params, want_files=True):
        buf = np.frombuffer(text, np.uint8)
        cap = max(1, len(buf) // 2)
        bwt = np.empty(cap, np.uint8) if want_files else None
        qs = np.empty(cap, np.uint8) if want_files else None
        lcp = np.empty(cap, np.uint8) if want_files else None
        out = np.empty(len(buf) + len(buf) // 4 + 64, np.uint8)
        size = C.c_uint64(0); nreads = C.c_uint64(0); n = C.c_uint64(0); st = Stats()
        self._chk(self.lib.bfq_pipeline_fastq(self.h, _ptr(buf), C.c_uint64(len(buf)), C.byref(params), _ptr(bwt), _ptr(qs), _ptr(lcp),
                                              C.c_uint64(cap), _ptr(out), C.c_uint64(len(out)), C.byref(size), C.byref(nreads),
                                              C.byref(n), C.byref(st)))
        k = int(n.value)
        files = (bwt[:k], qs[:k], lcp[:k]) if want_files else None
        return files, out[:int(size.value)].tobytes(), st.as_dict()

    # ---- quality-stream pre-transform for the host entropy coders
    def qs_rle_encode(self, qs: bytes | np.ndarray):
        """The step-4 quality stream (one line per read) -> (run values, run lengths), two uint8 arrays of equal length."""
        buf = np.frombuffer(qs, np.uint8) if isinstance(qs, (bytes, bytearray)) else np.ascontiguousarray(qs, np.uint8)
        n = len(buf)
        vals = np.empty(max(n, 1), np.uint8); lens = np.empty(max(n, 1), np.uint8)
        nruns = C.c_uint64(0)
        self._chk(self.lib.bfq_qs_rle_encode(self.h, _ptr(buf), C.c_uint64(n), _ptr(vals), _ptr(lens), C.c_uint64(n), C.byref(nruns)))
        return vals[:int(nruns.value)], lens[:int(nruns.value)]

    def qs_rle_decode(self, vals, lens, capacity: int | None = None):
        vals = np.ascontiguousarray(vals, np.uint8); lens = np.ascontiguousarray(lens, np.uint8)
        cap = int(lens.astype(np.uint64).sum()) if capacity is None else capacity
        out = np.empty(max(cap, 1), np.uint8)
        n = C.c_uint64(0)
        self._chk(self.lib.bfq_qs_rle_decode(self.h, _ptr(vals), _ptr(lens), C.c_uint64(len(vals)), _ptr(out), C.c_uint64(cap), C.byref(n)))
        return out[:int(n.value)]

    # ---- one collection over several GPUs (bfq_dist_build / bfq_dist_place)
    def dist_build(self, d_bases, d_quals, d_off, nreads, total_bases, params: Params, rank, world, capacity,
                   text_bounds, d_bwt, d_qs, d_lcp, d_keys, d_vals):
        st = Stats()
        tb = np.ascontiguousarray(text_bounds, np.uint64)
        assert len(tb) == world + 1
        n_local = C.c_uint64(0); row_base = C.c_uint64(0)
        counts = np.zeros(world, np.uint64)
        self._chk(self.lib.bfq_dist_build(self.h, _ptr(d_bases), _ptr(d_quals), _ptr(d_off), C.c_uint64(nreads),
                                          C.c_uint64(total_bases), C.byref(params), C.c_int(rank), C.c_int(world),
                                          C.c_uint64(capacity), _ptr(tb), _ptr(d_bwt), _ptr(d_qs), _ptr(d_lcp),
                                          _ptr(d_keys), _ptr(d_vals), C.byref(n_local), C.byref(row_base),
                                          _ptr(counts), C.byref(st)))
        return int(n_local.value), int(row_base.value), counts, st.as_dict()

    def dist_place(self, d_keys, d_vals, count, text_lo, d_off, read_lo, read_hi, d_out_bases, d_out_quals):
        self._chk(self.lib.bfq_dist_place(self.h, _ptr(d_keys), _ptr(d_vals), C.c_uint64(count), C.c_uint64(text_lo),
                                          _ptr(d_off), C.c_uint64(read_lo), C.c_uint64(read_hi), _ptr(d_out_bases),
                                          _ptr(d_out_quals)))

    # ---- test hooks
    def debug_sort_pairs(self, keys, vals, bits):
        keys = np.ascontiguousarray(keys, np.uint64).copy(); vals = np.ascontiguousarray(vals, np.uint32).copy()
        self._chk(self.lib.bfq_debug_sort_pairs(self.h, _ptr(keys), _ptr(vals), C.c_uint64(len(keys)), C.c_int(bits)))
        return keys, vals

    def debug_suffix_array(self, bases, off):
        off = np.ascontiguousarray(off, np.uint64)
        nreads = len(off) - 1
        n = int(off[-1]) + nreads
        sa = np.empty(n, np.uint32)
        self._chk(self.lib.bfq_debug_suffix_array(self.h, _ptr(bases), _ptr(off), C.c_uint64(nreads), _ptr(sa)))
        return sa


NOALLOC_MAX_PAIRS = 1 << 22          # BFQ_NOALLOC_MAX_PAIRS (include/bfq_gpu.h)


def sort_pairs_noalloc(d_keys_a, d_keys_b, d_vals_a, d_vals_b, n: int, bits: int, cuda_stream: int = 0, M: int = 2, B: int = 0) -> int:
    """bfq_sort_pairs_noalloc: context-free, allocation-free radix sort of device-resident (u64, u32) pairs on the
    current device; asynchronous on ``cuda_stream``.  Returns 0 when the result is in the (a) buffers, 1 for (b)."""
    lib = load(M, B)
    flag = C.c_int(0)
    rc = lib.bfq_sort_pairs_noalloc(_ptr(d_keys_a), _ptr(d_keys_b), _ptr(d_vals_a), _ptr(d_vals_b), C.c_uint64(n),
                                    C.c_int(bits), C.c_void_p(cuda_stream or 0), C.byref(flag))
    if rc != 0:
        raise BfqError(rc, lib.bfq_last_error().decode())
    return int(flag.value)
