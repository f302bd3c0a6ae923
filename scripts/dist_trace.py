#!/usr/bin/env python
"""Stage timeline (CUDA events at stage boundaries, option trace=1) of ONE rank's build in single-collection mode on one GPU:
rank `--rank` of `--world` on a collection of --nreads x --len, 64-bit positions forced with --wide.  Shows where a rank's
time goes when the replicated text is large and its own slice small (the N=8 case of BASELINE config 5).

    python scripts/dist_trace.py --nreads 25000000 --len 150 --world 8 --rank 3 --wide
"""
import argparse, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from bfqzip_b200 import dist as bdist, lib, synth

ap = argparse.ArgumentParser()
ap.add_argument("--nreads", type=int, default=25_000_000); ap.add_argument("--len", type=int, default=150)
ap.add_argument("--world", type=int, default=8); ap.add_argument("--rank", type=int, default=3)
ap.add_argument("--wide", action="store_true"); ap.add_argument("--steps", type=int, default=2)
a = ap.parse_args()
dev = torch.device("cuda", 0)
glen = synth.hash_genome_len(a.nreads, a.len, 40.0, a.len)
d_b, d_q = synth.hash_reads_torch(torch, dev, a.nreads, a.len, 1004, glen)
off = np.arange(a.nreads + 1, dtype=np.uint64) * np.uint64(a.len)
dreads = (d_b, d_q, torch.from_numpy(off.view(np.int64)).to(dev))
torch.cuda.synchronize(); torch.cuda.empty_cache()
eng = bdist.GpuEngine(0, 0, 0, {"wide_positions": 1} if a.wide else None)
eng.g.set_option("profile", 1)
rb = bdist.read_bounds(off, a.world); tb = bdist.text_bounds(off, rb)
P = lib.default_params(eos=0)
for s in range(a.steps):
    if s == a.steps - 1:
        eng.g.set_option("trace", 1)
    eng.g.prof_reset()
    t0 = time.perf_counter()
    sl = eng.build(a.rank, a.world, dreads, a.nreads, int(off[-1]), P, tb)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) * 1e3
    cls = {k: round(eng.g.prof_get(k)[0], 2) for k in lib.PROF_CLASSES if eng.g.prof_get(k)[0] > 0}
    print(f"step {s}: build {dt:.1f} ms, slice {sl['n_local']} rows of {int(off[-1]) + a.nreads}; kernel classes sum {sum(cls.values()):.1f} ms: {cls}", flush=True)
