#!/usr/bin/env python
"""A/B of one context option on the device-resident pipeline: same input, option = each given value in turn; the five
outputs must be bit-identical between the runs; prints the stage times of the last of --steps runs per value.

    python scripts/ab_option.py bulk_tile 0 1 [--cfg4]      (--cfg4: one config-4 block, 2 x 2.5M x 150 bp, M=3)
"""
import hashlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from bfqzip_b200 import lib, synth

args = [a for a in sys.argv[1:] if not a.startswith("--")]
opt, values = args[0], [int(v) for v in args[1:]]
cfg4 = "--cfg4" in sys.argv
steps = 4
dev = torch.device("cuda", 0)
if cfg4:
    npairs, L, M = 20_000_000, 150, 3
    g = synth.hash_genome_len(2 * npairs, L, 40.0, 350)
    d_b, d_q, off = synth.hash_paired_block_torch(torch, dev, npairs, L, 1003, g, 350, 0, npairs // 8)
else:
    M = 2
    b, q, off = synth.make_reads(1_000_000, 100, 1001)
    d_b, d_q = torch.from_numpy(b).to(dev), torch.from_numpy(q).to(dev)
nreads = len(off) - 1; total = int(off[-1]); N = total + nreads
d_off = torch.from_numpy(off.view(np.int64)).to(dev)
u8 = lambda k: torch.empty(k + 64, dtype=torch.uint8, device=dev)
outs = [u8(N), u8(N), u8(N), u8(total), u8(total)]
torch.cuda.synchronize(); torch.cuda.empty_cache()
md5 = lambda t, k: hashlib.md5(t[:k].cpu().numpy().tobytes()).hexdigest()
ref = None
for v in values:
    c = lib.BfqGpu(M, 0, 0)
    c.set_option("profile", 1); c.set_option(opt, v)
    for s in range(steps):
        c.prof_reset()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        st = c.pipeline_dev(d_b, d_q, d_off, nreads, total, lib.default_params(eos=0), *outs)
        e1.record(); torch.cuda.synchronize()
    stages = {k: round(c.prof_get(k)[0], 3) for k in lib.PROF_CLASSES if c.prof_get(k)[0] > 0}
    dig = [md5(outs[0], N), md5(outs[1], N), md5(outs[2], N), md5(outs[3], total), md5(outs[4], total)]
    same = ref is None or dig == ref
    ref = ref or dig
    print(json.dumps({opt: v, "ms": round(e0.elapsed_time(e1), 3), "identical_to_first": same, "stages_ms": stages}), flush=True)
    c.close()
    if not same:
        raise SystemExit(f"outputs differ with {opt}={v}")
