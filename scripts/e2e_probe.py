#!/usr/bin/env python
"""Where the host-buffer call spends its time: full call, call without the three files, device-only, and the copies alone."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from bfqzip_b200 import lib, synth
bases, quals, off = synth.make_reads(1_000_000, 100, 1001)
nreads = len(off) - 1; total = int(off[-1]); N = total + nreads
dev = torch.device("cuda", 0)
g = lib.BfqGpu(2, 0, 0); P = lib.default_params(eos=0)
for kv in sys.argv[1:]:
    k, v = kv.split("="); g.set_option(k, int(v))
pin = lambda a: torch.from_numpy(a).pin_memory()
h_b, h_q, h_off = pin(bases), pin(quals), pin(off.view(np.int64))
e = lambda n: torch.empty(n, dtype=torch.uint8).pin_memory()
h_bwt, h_qs, h_lcp, h_ob, h_oq = e(N), e(N), e(N), e(total), e(total)
d_b = torch.empty(total, dtype=torch.uint8, device=dev); d_q = torch.empty_like(d_b); d_off = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
de = lambda n: torch.empty(n + 64, dtype=torch.uint8, device=dev)
d_bwt, d_qs, d_lcp, d_ob, d_oq = de(N), de(N), de(N), de(total), de(total)
def full(): g.pipeline(h_b, h_q, h_off, P, out=(h_bwt, h_qs, h_lcp, h_ob.numpy(), h_oq.numpy()))
def nofiles(): g.pipeline(h_b, h_q, h_off, P, out=(None, None, None, h_ob.numpy(), h_oq.numpy()))
def h2d():
    d_b.copy_(h_b, non_blocking=True); d_q.copy_(h_q, non_blocking=True); d_off.copy_(h_off, non_blocking=True); torch.cuda.synchronize()
def devonly(): g.pipeline_dev(d_b, d_q, d_off, nreads, total, P, d_bwt, d_qs, d_lcp, d_ob, d_oq)
def d2h():
    h_bwt.copy_(d_bwt[:N], non_blocking=True); h_qs.copy_(d_qs[:N], non_blocking=True); h_lcp.copy_(d_lcp[:N], non_blocking=True)
    h_ob.copy_(d_ob[:total], non_blocking=True); h_oq.copy_(d_oq[:total], non_blocking=True); torch.cuda.synchronize()
def serial(): h2d(); devonly(); d2h()
def wall(fn, reps=5):
    fn(); fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return round((time.perf_counter() - t0) / reps * 1e3, 3)
print(json.dumps({k: wall(f) for k, f in (("full", full), ("nofiles", nofiles), ("h2d", h2d), ("devonly", devonly), ("d2h", d2h), ("serial", serial))}))
