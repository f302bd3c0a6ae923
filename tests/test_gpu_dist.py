"""Single-collection mode over several GPUs (bfq_dist_build / bfq_dist_place, bfqzip_b200/dist.py): the ranks'
slices and rebuilt read ranges, put together, must be bit-identical to the oracle on the whole collection.
On one GPU the ranks run one after the other (dist.emulate_ranks); the 2-GPU run is scripts/dist_check.py."""
import os

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import dist, lib, synth

pytestmark = pytest.mark.gpu
STAT_KEYS = ("clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified", "nreads")


def check(reads, M, B, world, eos=0, K=16, options=None):
    bases, quals, off = reads
    eng = dist.GpuEngine(M, B, 0, options)
    try:
        bwt, qs, lcp, ob, oq, st = dist.emulate_ranks(eng, reads, lib.default_params(eos=eos, K=K), world)
    finally:
        eng.close()
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, eos)
    assert np.array_equal(bwt, rb) and np.array_equal(qs, rq)
    assert np.array_equal(lcp, rl)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B, eos=eos, K=K))
    assert np.array_equal(ob, eb) and np.array_equal(oq, eq)
    assert all(st[k] == est[k] for k in STAT_KEYS)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_dist_matches_oracle(world):
    check(synth.make_reads(20000, 100, 900 + world), 2, 0, world)


@pytest.mark.parametrize("world", [1, 2, 5])
def test_dist_wide_positions(world):
    """64-bit text positions (what collections of 2^32 rows and more use), forced on a small collection."""
    check(synth.make_reads(15000, 100, 31 + world), 2, 0, world, options={"wide_positions": 1})
    check(synth.make_reads(4000, 70, 32 + world), 0, 1, world, options={"wide_positions": 1, "group_cap": 8})


@pytest.mark.parametrize("mode", [(0, 1), (1, 0), (3, 1)], ids=lambda m: f"M{m[0]}B{m[1]}")
def test_dist_modes(mode):
    check(synth.make_reads(12000, 80, 77), mode[0], mode[1], 3)


def test_dist_variable_length_and_duplicates():
    g = np.load(os.path.join(H.GOLDEN, "syn_var_dup.npz"))
    check((g["bases"], g["quals"], g["off"]), 3, 1, 4, eos=35)


def test_dist_small_k_and_deferred_groups():
    # -k 4 is the smallest context the mode accepts; a tiny group cap forces the deferred levels inside every slice
    reads = synth.make_reads(6000, 60, 5)
    check(reads, 2, 0, 2, K=4)
    check(reads, 2, 0, 3, options={"group_cap": 8})


def test_dist_more_ranks_than_work():
    rng = np.random.default_rng(3)
    bases = rng.choice(np.frombuffer(b"ACGT", np.uint8), 5 * 30).astype(np.uint8)
    quals = rng.integers(40, 70, 150).astype(np.uint8)
    off = (np.arange(6) * 30).astype(np.uint64)
    check((bases, quals, off), 2, 0, 8)


def test_dist_refuses_k_below_4():
    reads = synth.make_reads(1000, 50, 1)
    eng = dist.GpuEngine(2, 0, 0)
    try:
        with pytest.raises(lib.BfqError):
            dist.emulate_ranks(eng, reads, lib.default_params(eos=0, K=3), 2)
    finally:
        eng.close()


def test_dist_place_refuses_a_pair_list_that_is_not_a_permutation():
    """bfq_dist_place is a public entry point: its keys come from an exchange outside the library.  A position outside the
    rank's range, a doubled position or a range that does not match the reads must be refused (BFQ_EINVAL), not scattered."""
    import torch
    reads = synth.make_reads(3000, 80, 11)
    off = np.ascontiguousarray(reads[2], np.uint64)
    eng = dist.GpuEngine(2, 0, 0)
    try:
        nreads, total = len(off) - 1, int(off[-1])
        tb = dist.text_bounds(off, [0, nreads])
        dreads = eng.upload(reads)
        sl = eng.build(0, 1, dreads, nreads, total, lib.default_params(eos=0), tb)
        cnt = int(sl["counts"][0])
        good_k, good_v = sl["keys"].clone(), sl["vals"].clone()
        torch.cuda.synchronize()
        ob, oq = eng.place(good_k.clone(), good_v.clone(), cnt, 0, dreads, 0, nreads, total)      # the intact list places
        assert ob.numel() == total
        for what in ("out_of_range", "doubled", "wrong_count", "wrong_start"):
            k, v = good_k.clone(), good_v.clone()
            args = dict(count=cnt, text_lo=0, read_lo=0, read_hi=nreads)
            if what == "out_of_range":
                k[cnt // 2] = cnt + 5
            elif what == "doubled":
                k[cnt // 3] = k[cnt // 3 + 1]
            elif what == "wrong_count":
                args["count"] = cnt - 16
            else:
                args["text_lo"] = 7
            torch.cuda.synchronize()
            with pytest.raises(lib.BfqError) as ei:
                eng.place(k, v, args["count"], args["text_lo"], dreads, args["read_lo"], args["read_hi"], total)
            assert ei.value.code == -3, what
        ob2, oq2 = eng.place(good_k.clone(), good_v.clone(), cnt, 0, dreads, 0, nreads, total)     # and the context still works
        assert torch.equal(ob, ob2) and torch.equal(oq, oq2)
    finally:
        eng.close()
