"""the quality-stream pre-transform for the host entropy coders (bfq_qs_rle_encode / _decode,
`bfq_ext -R`): run values + run lengths of the step-4 stream OUT.fq.qs.  Checked against the numpy restatement
(helpers.rle_encode_numpy) and through the round trip; there is no counterpart in the reference."""
import os
import subprocess

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import lib, synth

pytestmark = pytest.mark.gpu
BIN = os.path.join(H.ROOT, "tools", "bin")


def qs_stream(quals, off):
    """The stream step 4 makes (sed -n 4~4p): every read's qualities followed by a newline."""
    o = off.astype(np.int64)
    nreads = len(o) - 1
    out = np.empty(len(quals) + nreads, np.uint8)
    idx = np.arange(len(quals), dtype=np.int64) + np.repeat(np.arange(nreads, dtype=np.int64), np.diff(o))
    out[idx] = quals
    out[o[1:] + np.arange(nreads, dtype=np.int64)] = 10
    return out


@pytest.mark.parametrize("M", [0, 2])
def test_rle_of_smoothed_stream_matches_numpy_and_round_trips(M):
    bases, quals, off = synth.make_reads(20000, 100, 77 + M, var_len=True)
    g = lib.BfqGpu(M, 0, 0)
    try:
        _, (ob, oq), st = g.pipeline(bases, quals, off, lib.default_params(eos=0), want_files=False)
        raw, smooth = qs_stream(quals, off), qs_stream(oq, off)
        for s in (raw, smooth):
            v, l = g.qs_rle_encode(s)
            ev, el = H.rle_encode_numpy(s)
            assert np.array_equal(v, ev) and np.array_equal(l, el)
            assert np.array_equal(g.qs_rle_decode(v, l), s)
        # what the transform is for: smoothing leaves long constant runs
        # (M=2 writes one constant; M=0 writes each cluster's maximum and leaves the other rows' qualities: fewer equal runs)
        assert len(g.qs_rle_encode(smooth)[0]) < {0: 0.8, 2: 0.6}[M] * len(g.qs_rle_encode(raw)[0])
        assert st["qs_smoothed"] > 0
    finally:
        g.close()


def test_rle_edge_cases():
    g = lib.BfqGpu(2, 0, 0)
    try:
        assert len(g.qs_rle_encode(b"")[0]) == 0 and len(g.qs_rle_decode(np.empty(0, np.uint8), np.empty(0, np.uint8))) == 0
        for s in (b"A", b"\n", b"AB", b"AA\n", b"I" * 255 + b"\n" + b"I" * 255 + b"\n", b"ABABABAB" * 1000,
                  bytes(np.random.default_rng(1).integers(33, 36, 100001, dtype=np.uint8))):
            a = np.frombuffer(s, np.uint8)
            if a.size > 1000:          # keep the runs of the random stream within a line's length
                a = a.copy(); a[::200] = 10
            v, l = g.qs_rle_encode(a)
            ev, el = H.rle_encode_numpy(a)
            assert np.array_equal(v, ev) and np.array_equal(l, el), s[:20]
            assert np.array_equal(g.qs_rle_decode(v, l), a)
        # lengths that cross the 16-byte and 4096-byte tile boundaries of the kernels
        rng = np.random.default_rng(2)
        lens = rng.integers(1, 256, 5000)
        vals = np.arange(5000) % 90 + 33
        a = np.repeat(vals.astype(np.uint8), lens)
        v, l = g.qs_rle_encode(a)
        assert np.array_equal(v, vals.astype(np.uint8)) and np.array_equal(l, lens.astype(np.uint8))
        # a run longer than 255 is outside the transform's domain
        with pytest.raises(lib.BfqError) as ei:
            g.qs_rle_encode(b"I" * 300)
        assert ei.value.code == -3 and "255" in str(ei.value)
        with pytest.raises(lib.BfqError) as ei:
            g.qs_rle_decode(np.array([65], np.uint8), np.array([9], np.uint8), capacity=4)
        assert ei.value.code == -8
    finally:
        g.close()


def test_rle_round_trip_at_config2_size():
    """1M x 100 bp: the whole smoothed stream of BASELINE config 2 (101 MB), round trip + run statistics."""
    bases, quals, off = synth.make_reads(1_000_000, 100, 1001)
    g = lib.BfqGpu(2, 0, 0)
    try:
        _, (ob, oq), st = g.pipeline(bases, quals, off, lib.default_params(eos=0), want_files=False)
        s = qs_stream(oq, off)
        v, l = g.qs_rle_encode(s)
        assert int(l.astype(np.int64).sum()) == len(s) and int(l.min()) >= 1
        assert not (v[1:] == v[:-1]).any()                    # runs are maximal
        assert np.array_equal(g.qs_rle_decode(v, l), s)
        assert len(v) < 0.45 * len(s)
    finally:
        g.close()


def test_bfq_ext_R_writes_the_two_streams(tmp_path):
    g0 = np.load(os.path.join(H.GOLDEN, "example.npz"))
    for e, a in ((".bwt", g0["bwt"]), (".bwt.qs", g0["qs"]), (".1.lcp", g0["lcp"])):
        a.tofile(str(tmp_path / ("o" + e)))
    r = subprocess.run([os.path.join(BIN, "bfq_ext"), "-e", "o.bwt", "-q", "o.bwt.qs", "-a", "o.1.lcp", "-o", "o", "-l", "250", "-s", "35", "-m", "5", "-R"],
                       cwd=tmp_path, capture_output=True)
    assert r.returncode == 0, r.stderr.decode()[-500:]
    qs = np.fromfile(tmp_path / "o.fq.qs", np.uint8)
    v, l = np.fromfile(tmp_path / "o.fq.qs.rlv", np.uint8), np.fromfile(tmp_path / "o.fq.qs.rll", np.uint8)
    ev, el = H.rle_encode_numpy(qs)
    assert np.array_equal(v, ev) and np.array_equal(l, el)
    assert np.array_equal(np.repeat(v, l), qs)
    assert (tmp_path / "o.fq.dna").exists()
