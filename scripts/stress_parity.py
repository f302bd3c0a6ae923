#!/usr/bin/env python
"""Randomised parity sweep on a GPU box: the fused pipeline against the oracle over coverages, read lengths, error
rates, variable lengths, modes and the group-ordering options (batched / per-group, flat / walk).  Test infrastructure:
it calls the oracle as the checker, like tests/ do.

    python scripts/stress_parity.py [--cases 24] [--seed 1]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def dist_sweep(a):
    """Single-collection mode (bfq_dist_build / bfq_dist_place): W emulated ranks on one GPU against the oracle on the whole collection."""
    import helpers as H
    from bfqzip_b200 import dist, lib, synth
    rng = np.random.default_rng(a.seed)
    bad = 0
    for case in range(a.cases):
        M, B = int(rng.integers(0, 4)), int(rng.integers(0, 2))
        W = int(rng.choice([1, 2, 3, 4, 5, 8]))
        cov = float(rng.choice([3, 8, 40, 70])); L = int(rng.choice([30, 75, 100, 150, 250]))
        nreads = int(rng.integers(1500, 7000)); err = float(rng.choice([0.0, 0.01, 0.05])); var_len = bool(rng.integers(0, 2))
        K, m = int(rng.choice([4, 8, 16, 16, 30])), int(rng.choice([2, 3, 5, 5, 8]))          # the mode needs -k >= 4
        qthr, fthr = int(rng.choice([10, 20, 20, 35])), float(rng.choice([34, 40, 40, 50, 60]))
        opts = {"wide_positions": int(rng.integers(0, 2)), "group_cap": int(rng.choice([8, 1 << 20])), "bulk_tile": int(rng.integers(0, 2))}
        reads = synth.make_reads(nreads, L, 5000 + case, coverage=cov, err=err, var_len=var_len)
        eng = dist.GpuEngine(M, B, 0, opts)
        try:
            bwt, qs, lcp, ob, oq, st = dist.emulate_ranks(eng, reads, lib.default_params(eos=0, K=K, m=m, qthr=qthr, fthr=fthr), W)
        finally:
            eng.close()
        rb, rq, rl = H.oracle_build_ebwt(*reads, 0)
        ok_a = np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)
        eb, eq, eoff, est, _, _, e = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B, K=K, m=m, qthr=qthr, fthr=fthr, eos=0))
        ok_c = e == 0 and np.array_equal(ob, eb) and np.array_equal(oq, eq) and all(
            st[k] == est[k] for k in ("clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"))
        print(f"dist case {case}: W={W} M={M} B={B} K={K} m={m} t={qthr} f={fthr} cov={cov} L={L} reads={nreads} err={err} var_len={var_len} {opts} "
              f"partA={'ok' if ok_a else 'DIFF'} partBC={'ok' if ok_c else 'DIFF'}", flush=True)
        bad += (not ok_a) + (not ok_c)
    print("stress_parity --dist:", "all ok" if bad == 0 else f"{bad} differences")
    sys.exit(1 if bad else 0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=24)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--big", action="store_true", help="30 000 - 120 000 reads per case instead of 2 000 - 9 000 (the oracle needs up to a minute per case)")
    ap.add_argument("--dist", action="store_true", help="the single-collection mode instead: dist.emulate_ranks with a random number of ranks")
    a = ap.parse_args()
    if a.dist:
        return dist_sweep(a)
    import helpers as H
    from bfqzip_b200 import lib, synth

    rng = np.random.default_rng(a.seed)
    ctxs = {}
    bad = 0
    for case in range(a.cases):
        M, B = int(rng.integers(0, 4)), int(rng.integers(0, 2))
        cov = float(rng.choice([2, 3, 5, 8, 15, 40, 70]))
        L = int(rng.choice([30, 75, 100, 150, 250]))
        nreads = int(rng.integers(30000, 120000)) if a.big else int(rng.integers(2000, 9000))
        err = float(rng.choice([0.0, 0.01, 0.05]))
        var_len = bool(rng.integers(0, 2))
        opts = {"batch_min_groups": int(rng.choice([0, 64, 1 << 20])), "rebuild": int(rng.choice([0, 1, 2])),
                "long_cluster": int(rng.choice([4, 16, 64, 1024])), "bulk_tile": int(rng.integers(0, 2))}
        # the cluster parameters of bfq_ext (-k -m -t -f -v); f >= 34 keeps at most two frequent symbols (the reference asserts otherwise)
        K, m = int(rng.choice([1, 2, 4, 8, 16, 16, 30])), int(rng.choice([2, 3, 5, 5, 8]))
        qthr, fthr, dq = int(rng.choice([10, 20, 20, 35])), float(rng.choice([34, 40, 40, 50, 60])), int(rng.choice([62, 35, 73]))
        bases, quals, off = synth.make_reads(nreads, L, 1000 + case, coverage=cov, err=err, var_len=var_len)
        if (M, B) not in ctxs:
            ctxs[(M, B)] = lib.BfqGpu(M, B, 0)
        g = ctxs[(M, B)]
        for k, v in opts.items():
            g.set_option(k, v)
        P = lib.default_params(eos=35, K=K, m=m, qthr=qthr, fthr=fthr, default_q=dq)
        (bwt, qs, lcp), (ob, oq), st = g.pipeline(bases, quals, off, P)
        rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
        ok_a = np.array_equal(bwt, rb) and np.array_equal(qs, rq) and np.array_equal(lcp, rl)
        eb, eq, eoff, est, _, _, e = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B, K=K, m=m, qthr=qthr, fthr=fthr, default_q=dq))
        ok_c = e == 0 and np.array_equal(ob, eb) and np.array_equal(oq, eq) and all(
            st[k] == est[k] for k in ("clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"))
        print(f"case {case}: M={M} B={B} K={K} m={m} t={qthr} f={fthr} v={dq} cov={cov} L={L} reads={nreads} err={err} var_len={var_len} {opts} "
              f"partA={'ok' if ok_a else 'DIFF'} partBC={'ok' if ok_c else 'DIFF'}", flush=True)
        bad += (not ok_a) + (not ok_c)
    print("stress_parity:", "all ok" if bad == 0 else f"{bad} differences")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
