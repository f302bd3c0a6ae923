"""GPU parity: the CUDA path through the C ABI vs the oracle / golden vectors (bit-exact)."""
import glob
import os

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import lib, synth

pytestmark = pytest.mark.gpu

MODES = [(M, B) for M in range(4) for B in range(2)]
STAT_KEYS = ["clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"]
GOLD = sorted(glob.glob(os.path.join(H.GOLDEN, "*.npz")))
GIDS = [os.path.basename(p)[:-4] for p in GOLD]

_ctx = {}


def ctx(M=2, B=0):
    if (M, B) not in _ctx:
        _ctx[(M, B)] = lib.BfqGpu(M, B, 0)
    return _ctx[(M, B)]


# ---------------------------------------------------------------- radix sort
@pytest.mark.parametrize("n,bits", [(1, 8), (5, 3), (255, 8), (4096, 16), (4097, 13), (100000, 48),
                                    (1 << 20, 54), (3000001, 64), (70000, 1)])
def test_radix_sort_pairs(n, bits):
    rng = np.random.default_rng(n + bits)
    keys = rng.integers(0, 1 << 63, size=n, dtype=np.uint64)
    if bits < 64:
        keys &= np.uint64((1 << bits) - 1)
    if n > 1000:            # many duplicates: stability matters
        keys[: n // 2] = keys[rng.integers(0, 50, size=n // 2)]
    vals = np.arange(n, dtype=np.uint32)
    k2, v2 = ctx().debug_sort_pairs(keys, vals, bits)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k2, keys[order])
    assert np.array_equal(v2, vals[order])


def test_radix_sort_ignores_high_bits():
    rng = np.random.default_rng(5)
    keys = rng.integers(0, 1 << 40, size=50000, dtype=np.uint64)
    vals = np.arange(50000, dtype=np.uint32)
    k2, v2 = ctx().debug_sort_pairs(keys, vals, 16)
    order = np.argsort(keys & np.uint64(0xFFFF), kind="stable")
    assert np.array_equal(v2, vals[order])


# ---------------------------------------------------------------- part A
@pytest.mark.parametrize("path", GOLD, ids=GIDS)
def test_part_a_golden(path):
    g = np.load(path)
    bwt, qs, lcp = ctx().build_ebwt(g["bases"], g["quals"], g["off"], 35)
    assert np.array_equal(bwt, g["bwt"])
    assert np.array_equal(qs, g["qs"])
    assert np.array_equal(lcp, g["lcp"])


@pytest.mark.parametrize("k", [1, 3, 7, 10, 16, 21])
def test_part_a_initial_k(k):
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    c = ctx()
    c.set_option("initial_k", k)
    try:
        bwt, qs, lcp = c.build_ebwt(g["bases"], g["quals"], g["off"], 35)
    finally:
        c.set_option("initial_k", 0)
    assert np.array_equal(bwt, g["bwt"]) and np.array_equal(qs, g["qs"]) and np.array_equal(lcp, g["lcp"])


@pytest.mark.parametrize("cap,k", [(1, 16), (3, 16), (40, 5), (1, 2), (300, 1)])
@pytest.mark.parametrize("name", ["syn_var_dup", "syn_short"])
def test_part_a_msd_levels(name, cap, k):
    """Small group_cap / short keys force the deferred-group (next MSD level) path on every group."""
    g = np.load(os.path.join(H.GOLDEN, name + ".npz"))
    c = ctx()
    c.set_option("group_cap", cap); c.set_option("initial_k", k)
    try:
        bwt, qs, lcp = c.build_ebwt(g["bases"], g["quals"], g["off"], 35)
        sa = c.debug_suffix_array(g["bases"], g["off"])
    finally:
        c.set_option("group_cap", 1 << 20); c.set_option("initial_k", 0)
    assert np.array_equal(bwt, g["bwt"]) and np.array_equal(qs, g["qs"]) and np.array_equal(lcp, g["lcp"])
    assert np.array_equal(np.sort(sa), np.arange(len(sa), dtype=np.uint32))


def test_part_a_huge_groups():
    """Repeats: groups far larger than a tile (poly-A reads, tandem repeats, many identical reads)."""
    rng = np.random.default_rng(9)
    reads = [b"A" * 120] * 900 + [b"ACACACACAC" * 9] * 700 + [b"ACGTTGCA" * 12 + b"N"] * 500
    reads += [bytes(rng.choice(list(b"ACGT"), size=int(rng.integers(20, 130))).astype(np.uint8)) for _ in range(800)]
    order = rng.permutation(len(reads))
    reads = [reads[i] for i in order]
    bases = np.frombuffer(b"".join(reads), dtype=np.uint8)
    off = np.zeros(len(reads) + 1, np.uint64); off[1:] = np.cumsum([len(r) for r in reads])
    quals = rng.integers(35, 74, size=bases.size, dtype=np.uint8)
    bwt, qs, lcp = ctx().build_ebwt(bases, quals, off, 35)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_part_a_majority_reference(seed):
    """In-group ordering compares every member with a string voted from the three longest members: groups whose longest
    members disagree pairwise at the same symbol (the vote yields a symbol none of them has), whose longest member carries
    errors, with two or three members only, with many members of equal length, and with N inside the tails."""
    rng = np.random.default_rng(seed)
    L = 90
    loci = [bytes(rng.choice(list(b"ACGT"), size=2 * L).astype(np.uint8)) for _ in range(40)]
    reads = []
    for li, g in enumerate(loci):
        cover = [2, 3, 5, 31, 33, 47, 64, 70][li % 8]             # group sizes around the 32 / 64 row paths
        for r in range(cover):
            s0 = int(rng.integers(0, L // 2))
            n = int(rng.integers(30, L)) if li % 3 else L             # li % 3 == 0: every member has the same length
            rd = bytearray(g[s0:s0 + n])
            mode = (li + r) % 5
            if mode == 0 and len(rd) > 40:                            # all reads of the locus disagree at one fixed symbol
                rd[40 - s0 if 40 - s0 >= 0 else 0] = b"ACGTN"[r % 5]
            elif mode == 1:                                           # a random error
                i = int(rng.integers(0, len(rd))); rd[i] = b"ACGT"[int(rng.integers(0, 4))]
            elif mode == 2 and len(rd) > 25:
                rd[int(rng.integers(20, len(rd)))] = ord("N")
            reads.append(bytes(rd))
    reads += [bytes(loci[0][:70])] * 5                                # identical reads: ties go to read order
    order = rng.permutation(len(reads))
    reads = [reads[i] for i in order]
    bases = np.frombuffer(b"".join(reads), dtype=np.uint8)
    off = np.zeros(len(reads) + 1, np.uint64); off[1:] = np.cumsum([len(r) for r in reads])
    quals = rng.integers(35, 74, size=bases.size, dtype=np.uint8)
    c = ctx()
    for k in (0, 6):                                                  # default key and a short one (larger groups)
        c.set_option("initial_k", k)
        try:
            bwt, qs, lcp = c.build_ebwt(bases, quals, off, 35)
        finally:
            c.set_option("initial_k", 0)
        rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
        assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)


@pytest.mark.parametrize("cov,min_groups", [(3, 0), (6, 0), (40, 0), (5, 64)])
def test_part_a_batched_small_groups(cov, min_groups):
    """Low coverage leaves a few rows per group: tiles with many groups order them in batches of whole groups and single
    rows (one warp per batch instead of one per group).  batch_min_groups = 0 forces that path on every tile."""
    nreads, L = 6000, 100
    bases, quals, off = synth.make_reads(nreads, L, 77 + cov, coverage=cov)
    c = ctx()
    c.set_option("batch_min_groups", min_groups)
    try:
        bwt, qs, lcp = c.build_ebwt(bases, quals, off, 35)
    finally:
        c.set_option("batch_min_groups", 64)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)


@pytest.mark.parametrize("cap", [2, 3, 5])
def test_part_a_batched_with_deferred_groups(cap):
    """Batches that contain groups handed to the next MSD level (group_cap): their rows stay put inside the batch."""
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    c = ctx()
    c.set_option("batch_min_groups", 0); c.set_option("group_cap", cap)
    try:
        bwt, qs, lcp = c.build_ebwt(g["bases"], g["quals"], g["off"], 35)
    finally:
        c.set_option("batch_min_groups", 64); c.set_option("group_cap", 1 << 20)
    assert np.array_equal(bwt, g["bwt"]) and np.array_equal(qs, g["qs"]) and np.array_equal(lcp, g["lcp"])


def test_part_a_refuses_reads_longer_than_255():
    """LCP entries are one byte and dataTypelenSeq is uchar (parameters.h:68): longer reads are an error, not a wrap."""
    bases = np.frombuffer(b"ACGT" * 80, dtype=np.uint8)
    quals = np.full(bases.size, 70, np.uint8)
    off = np.array([0, 300, 320], np.uint64)
    with pytest.raises(lib.BfqError) as ei:
        ctx().build_ebwt(bases, quals, off, 35)
    assert ei.value.code == -3


@pytest.mark.parametrize("eos", [0, 35])
def test_part_a_eos_byte(eos):
    """eGap writes 0, gsufsort '#': identical files otherwise."""
    g = np.load(os.path.join(H.GOLDEN, "example.npz"))
    bwt, qs, lcp = ctx().build_ebwt(g["bases"], g["quals"], g["off"], eos)
    ref = g["bwt"].copy(); ref[g["bwt"] == 35] = eos
    rq = g["qs"].copy(); rq[g["bwt"] == 35] = eos
    assert np.array_equal(bwt, ref) and np.array_equal(qs, rq) and np.array_equal(lcp, g["lcp"])


@pytest.mark.parametrize("seed,nreads,L,var", [(1, 20000, 100, False), (2, 30000, 150, True), (3, 5000, 250, True),
                                               (4, 1, 50, False), (5, 2, 1, False), (6, 60000, 36, False)])
def test_part_a_vs_oracle(seed, nreads, L, var):
    bases, quals, off = synth.make_reads(nreads, L, seed, var_len=var)
    bwt, qs, lcp = ctx().build_ebwt(bases, quals, off, 35)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)


def test_part_a_identical_reads():
    """All reads equal: every group is resolved only by the per-read end-markers."""
    one = np.frombuffer(b"ACGTNACGTTTGACA" * 6, dtype=np.uint8)
    R = 700
    bases = np.tile(one, R); quals = np.full(bases.size, 70, np.uint8)
    off = np.arange(R + 1, dtype=np.uint64) * np.uint64(one.size)
    bwt, qs, lcp = ctx().build_ebwt(bases, quals, off, 35)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)


# ---------------------------------------------------------------- parts B + C
@pytest.mark.parametrize("mode", MODES, ids=[f"M{m}B{b}" for m, b in MODES])
@pytest.mark.parametrize("path", GOLD, ids=GIDS)
def test_part_bc_golden(path, mode):
    M, B = mode
    g = np.load(path)
    P = lib.default_params(eos=35)
    ob, oq, off, st = ctx(M, B).smooth_rebuild(g["bwt"], g["qs"], g["lcp"], P)
    assert np.array_equal(off, g["off"])
    assert np.array_equal(ob, g[f"out_bases_M{M}_B{B}"])
    assert np.array_equal(oq, g[f"out_quals_M{M}_B{B}"])
    assert [st[k] for k in STAT_KEYS] == [int(x) for x in g[f"stats_M{M}_B{B}"]]


@pytest.mark.parametrize("mode", [(0, 0), (1, 0), (2, 0), (3, 0)], ids=lambda m: f"M{m[0]}")
def test_smoothed_rows_match_oracle(mode):
    """The per-row state (smoothed qualities, modified bases) before reconstruction."""
    M, B = mode
    g = np.load(os.path.join(H.GOLDEN, "syn_fixed.npz"))
    c = ctx(M, B)
    c.smooth_rebuild(g["bwt"], g["qs"], g["lcp"], lib.default_params(eos=35))
    q, b = c.export_rows(len(g["bwt"]))
    sq, ob, st, err = H.oracle_smooth(g["bwt"], g["qs"], g["lcp"], H.make_params(M=M, B=B))
    rows = g["bwt"] != 35
    assert np.array_equal(q[rows], sq[rows])
    assert np.array_equal(b, ob)


@pytest.mark.parametrize("long_len", [4, 16, 64])
@pytest.mark.parametrize("mode", [(0, 0), (1, 1), (2, 0), (3, 0)], ids=lambda m: f"M{m[0]}B{m[1]}")
def test_long_cluster_path(mode, long_len):
    """Force the block-per-cluster kernel and clusters that straddle tiles."""
    M, B = mode
    bases, quals, off = synth.make_reads(9000, 80, 77, coverage=200.0, err=0.03)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    c = ctx(M, B)
    c.set_option("long_cluster", long_len)
    try:
        ob, oq, o2, st = c.smooth_rebuild(rb, rq, rl, lib.default_params(eos=35))
    finally:
        c.set_option("long_cluster", 1024)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B))
    assert np.array_equal(o2, eoff) and np.array_equal(ob, eb) and np.array_equal(oq, eq)
    assert all(st[k] == est[k] for k in STAT_KEYS)


@pytest.mark.parametrize("K,m", [(16, 2), (8, 5), (30, 3), (1, 5), (200, 5)])
def test_params_K_m(K, m):
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    P = lib.default_params(eos=35, K=K, m=m)
    ob, oq, off, st = ctx(1, 0).smooth_rebuild(g["bwt"], g["qs"], g["lcp"], P)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(g["bwt"], g["qs"], g["lcp"], H.make_params(M=1, B=0, K=K, m=m))
    assert np.array_equal(ob, eb) and np.array_equal(oq, eq) and np.array_equal(off, eoff)
    assert all(st[k] == est[k] for k in STAT_KEYS)


# ---------------------------------------------------------------- fused pipeline
@pytest.mark.parametrize("rebuild", [1, 2], ids=["unpermute", "lfwalk"])
@pytest.mark.parametrize("mode", MODES, ids=[f"M{m}B{b}" for m, b in MODES])
def test_pipeline_vs_oracle(mode, rebuild):
    """Both part-C paths of the fused pipeline: streaming un-permute through the suffix array (flat mode: no
    walk lines, rule 11 reads the text) and the inverse-LF walk."""
    M, B = mode
    bases, quals, off = synth.make_reads(25000, 100, 500 + M * 2 + B)
    c = ctx(M, B)
    c.set_option("rebuild", rebuild)
    try:
        (bwt, qs, lcp), (ob, oq), st = c.pipeline(bases, quals, off, lib.default_params(eos=0))
        sq_g, ob_g = c.export_rows(len(bwt))
    finally:
        c.set_option("rebuild", 0)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 0)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B, eos=0))
    assert np.array_equal(ob, eb) and np.array_equal(oq, eq)
    assert all(st[k] == est[k] for k in STAT_KEYS)
    assert len(sq_g) == len(bwt) and len(ob_g) == len(bwt)


@pytest.mark.parametrize("rebuild", [1, 2], ids=["unpermute", "lfwalk"])
def test_pipeline_variable_length_and_duplicates(rebuild):
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    c = ctx(3, 1)
    c.set_option("rebuild", rebuild)
    try:
        (bwt, qs, lcp), (ob, oq), st = c.pipeline(g["bases"], g["quals"], g["off"], lib.default_params(eos=35))
    finally:
        c.set_option("rebuild", 0)
    assert np.array_equal(bwt, g["bwt"]) and np.array_equal(lcp, g["lcp"])
    assert np.array_equal(ob, g["out_bases_M3_B1"]) and np.array_equal(oq, g["out_quals_M3_B1"])
    assert [st[k] for k in STAT_KEYS] == [int(x) for x in g["stats_M3_B1"]]


def test_pipeline_roundtrip_1M():
    """BASELINE config 2 size (1M x 100 bp, M=2): size-independent properties at full size:
    bases may only change where the oracle rule allows (counted), qualities only inside clusters,
    read lengths are preserved, the eBWT is a permutation of the input symbols."""
    bases, quals, off = synth.make_reads(1_000_000, 100, 1001)
    (bwt, qs, lcp), (ob, oq), st = ctx(2, 0).pipeline(bases, quals, off, lib.default_params(eos=0))
    n = len(bwt)
    assert st["nreads"] == 1_000_000 and st["n"] == n
    assert np.array_equal(np.bincount(bwt, minlength=256)[1:], np.bincount(bases, minlength=256)[1:])
    assert int((bwt == 0).sum()) == 1_000_000
    assert np.array_equal(np.sort(qs[bwt != 0]), np.sort(quals))
    assert int((ob != bases).sum()) == st["modified"]
    changed = int((oq != quals).sum())
    assert changed <= st["qs_smoothed"]
    assert set(np.unique(oq[oq != quals]).tolist()) <= {ord(">")}
    assert lcp[0] == 0 and int(lcp.max()) <= 100
    # the host-buffer call hands finished rows / reads to the host slice by slice while later slices still run
    # (pageable and pinned destinations): it must return exactly what the device-buffer call leaves in HBM
    import torch
    dev = torch.device("cuda", 0)
    g = ctx(2, 0)
    d = lambda a: torch.from_numpy(a).to(dev)
    e = lambda k: torch.empty(k + 64, dtype=torch.uint8, device=dev)
    total = len(bases)
    d_bwt, d_qs, d_lcp, d_ob, d_oq = e(n), e(n), e(n), e(total), e(total)
    st2 = g.pipeline_dev(d(bases), d(quals), d(off.view(np.int64)), 1_000_000, total, lib.default_params(eos=0), d_bwt, d_qs, d_lcp, d_ob, d_oq)
    assert st2 == st
    for host, devt in ((bwt, d_bwt), (qs, d_qs), (lcp, d_lcp), (ob, d_ob), (oq, d_oq)):
        assert np.array_equal(host, devt[:len(host)].cpu().numpy())
    for slices in (1, 7):
        g.set_option("host_slices", slices)
        try:
            (b2, q2, l2), (ob2, oq2), st3 = g.pipeline(bases, quals, off, lib.default_params(eos=0))
        finally:
            g.set_option("host_slices", 16)
        assert st3 == st and np.array_equal(b2, bwt) and np.array_equal(q2, qs) and np.array_equal(l2, lcp)
        assert np.array_equal(ob2, ob) and np.array_equal(oq2, oq)


def test_fused_calls_refuse_reads_longer_than_max_len():
    """A read longer than -l is truncated by the reference (decode.cpp:605), which moves every later offset.  The fused
    calls return the reads under the CALLER's offsets, so they refuse such input (BFQ_EINVAL) instead of returning shifted
    reads; the two-call path truncates like the reference and returns the new offsets."""
    import torch
    bases, quals, off = synth.make_reads(300, 60, 4242, var_len=True)
    P = lib.default_params(eos=35, max_len=40)
    assert int(np.diff(off.astype(np.int64)).max()) > 40
    c = ctx(2, 0)
    with pytest.raises(lib.BfqError) as ei:
        c.pipeline(bases, quals, off, P)
    assert ei.value.code == -3 and "max_len" in str(ei.value)
    dev = torch.device("cuda", 0)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    n = len(bases) + len(off) - 1
    e = lambda k: torch.empty(k + 64, dtype=torch.uint8, device=dev)
    with pytest.raises(lib.BfqError) as ei:
        c.pipeline_dev(d(bases), d(quals), d(off.view(np.int64)), len(off) - 1, len(bases), P, e(n), e(n), e(n), e(n), e(n))
    assert ei.value.code == -3
    from bfqzip_b200 import fastq
    with pytest.raises(lib.BfqError) as ei:
        c.pipeline_fastq(fastq.format_fastq(bases, quals, off), P)
    assert ei.value.code == -3
    # the two-call path: truncated reads, new offsets, equal to the oracle's reconstruction with the same -l
    bwt, qs, lcp = c.build_ebwt(bases, quals, off, 35)
    ob, oq, o2, st = c.smooth_rebuild(bwt, qs, lcp, P)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(bwt, qs, lcp, H.make_params(M=2, B=0, max_len=40))
    assert np.array_equal(o2, eoff) and np.array_equal(ob, eb) and np.array_equal(oq, eq)
    assert int(np.diff(o2.astype(np.int64)).max()) == 40
    # and the context is usable afterwards
    (b2, q2, l2), _, _ = c.pipeline(bases, quals, off, lib.default_params(eos=35, max_len=250))
    assert np.array_equal(b2, bwt)


@pytest.mark.parametrize("bulk", [0, 1])
def test_part_a_tile_load_paths(bulk):
    """group_tile_kernel brings a window's keys and positions in either with two cp.async.bulk copies (mbarrier completion,
    the default) or with per-thread loads (the last, partial windows always do): both must give the reference's rows."""
    g = np.load(os.path.join(H.GOLDEN, "syn_fixed.npz"))
    bases, quals, off = synth.make_reads(30000, 100, 31337)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    c = ctx()
    c.set_option("bulk_tile", bulk)
    try:
        bwt, qs, lcp = c.build_ebwt(g["bases"], g["quals"], g["off"], 35)
        b2, q2, l2 = c.build_ebwt(bases, quals, off, 35)
    finally:
        c.set_option("bulk_tile", 1)
    assert np.array_equal(bwt, g["bwt"]) and np.array_equal(qs, g["qs"]) and np.array_equal(lcp, g["lcp"])
    assert np.array_equal(b2, rb) and np.array_equal(q2, rq) and np.array_equal(l2, rl)


def test_two_base_cluster_list_overflow_is_retried():
    """The list of two-frequent-base clusters (rule 11) is sized n / 8; a cluster needs only two rows, so valid input can
    hold more.  The kernels count what they would have stored and the smoothing pass is repeated once with lists of that
    size (r11_cap forces the overflow): same reads, same statistics as the oracle."""
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    P = lib.default_params(eos=35, K=8, m=2)
    c = ctx(3, 0)
    c.set_option("r11_cap", 1)
    try:
        ob, oq, off, st = c.smooth_rebuild(g["bwt"], g["qs"], g["lcp"], P)
    finally:
        c.set_option("r11_cap", 0)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(g["bwt"], g["qs"], g["lcp"], H.make_params(M=3, B=0, K=8, m=2))
    assert est["ambiguous"] + est["processed"] > 1            # the case really has more than one such cluster
    assert np.array_equal(ob, eb) and np.array_equal(oq, eq) and np.array_equal(off, eoff)
    assert all(st[k] == est[k] for k in STAT_KEYS)


@pytest.mark.parametrize("M", [1, 3])
def test_more_than_1024_clusters_per_tile(M):
    """-k 4 -m 3 at 2x coverage leaves ~1470 tiny clusters in every 4608-row window of cluster_tile_kernel: more than its
    1024 accumulators, so it works in rounds.  (Found by scripts/stress_parity.py: a row marked as replaced in one round
    made the packed quality compare of the next round count a neighbouring row as smoothed once too often.)"""
    bases, quals, off = synth.make_reads(3707, 250, 1123, coverage=2.0, err=0.05)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    P = lib.default_params(eos=35, K=4, m=3, qthr=10, fthr=40.0, default_q=73)
    ob, oq, o2, st = ctx(M, 0).smooth_rebuild(rb, rq, rl, P)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=0, K=4, m=3, qthr=10, fthr=40.0, default_q=73))
    assert np.array_equal(ob, eb) and np.array_equal(oq, eq) and np.array_equal(o2, eoff)
    assert {k: st[k] for k in STAT_KEYS} == {k: est[k] for k in STAT_KEYS}
