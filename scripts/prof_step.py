#!/usr/bin/env python
"""One or more fused pipeline steps on device-resident synthetic reads: the short command the
ncu captures under profiles/ are taken on (bench.py is too long to replay 40x per kernel).

    python scripts/prof_step.py [--nreads 1000000] [--len 100] [--M 2] [--B 0] [--steps 2] [--stages]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nreads", type=int, default=1_000_000)
    ap.add_argument("--len", type=int, default=100)
    ap.add_argument("--M", type=int, default=2)
    ap.add_argument("--B", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1001)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--opt", action="append", default=[], help="ctx option key=value (repeatable)")
    ap.add_argument("--stages", action="store_true", help="print the per-stage CUDA-event times")
    a = ap.parse_args()

    import torch
    from bfqzip_b200 import lib, synth
    bases, quals, off = synth.make_reads(a.nreads, a.len, a.seed)
    nreads = len(off) - 1
    total = int(off[-1]); N = total + nreads
    dev = torch.device("cuda", 0)
    g = lib.BfqGpu(a.M, a.B, 0)
    if a.stages:
        g.set_option("profile", 1)
    for kv in a.opt:
        k, v = kv.split("=")
        g.set_option(k, int(v))
    P = lib.default_params(eos=0)
    d_b = torch.from_numpy(bases).to(dev); d_q = torch.from_numpy(quals).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_bwt = torch.empty(N + 64, dtype=torch.uint8, device=dev); d_qs = torch.empty_like(d_bwt); d_lcp = torch.empty_like(d_bwt)
    d_ob = torch.empty(total + 64, dtype=torch.uint8, device=dev); d_oq = torch.empty_like(d_ob)
    torch.cuda.synchronize()
    import time
    for s in range(a.steps):
        g.prof_reset()
        t0 = time.perf_counter()
        st = g.pipeline_dev(d_b, d_q, d_off, nreads, total, P, d_bwt, d_qs, d_lcp, d_ob, d_oq)
        dt = time.perf_counter() - t0
        line = {"step": s, "ms": dt * 1e3, "gbases_s": total / dt / 1e9, "launches": g.launch_count()}
        if a.stages:
            line["stages_ms"] = {k: round(g.prof_get(k)[0], 3) for k in lib.PROF_CLASSES}
        print(json.dumps(line), flush=True)
    print(json.dumps(st))
    g.close()


if __name__ == "__main__":
    main()
