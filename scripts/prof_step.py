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


def torch_reads(torch, dev, nreads, L, seed, glen=0):
    """synth.make_reads' model (40x coverage, SNP haplotype, 1 % errors, an N every 1000th read) generated on the device."""
    gen = torch.Generator(device=dev); gen.manual_seed(seed)
    glen = glen or max(L + 1, nreads * L // 40)
    hap = torch.randint(0, 4, (2, glen), device=dev, dtype=torch.uint8, generator=gen)
    hap[1] = hap[0]
    snp = torch.arange(500, glen, 1000, device=dev)
    hap[1, snp] = (hap[0, snp] + torch.randint(1, 4, snp.shape, device=dev, dtype=torch.uint8, generator=gen)) & 3
    start = torch.randint(0, glen - L + 1, (nreads,), device=dev, generator=gen)
    which = torch.randint(0, 2, (nreads,), device=dev, generator=gen)
    acgt = torch.tensor(list(b"ACGT"), device=dev, dtype=torch.uint8)
    d_b = torch.empty((nreads, L), device=dev, dtype=torch.uint8); d_q = torch.empty_like(d_b)
    CH = 1 << 20
    for s0 in range(0, nreads, CH):
        s1 = min(nreads, s0 + CH)
        idx = start[s0:s1, None] + torch.arange(L, device=dev)[None, :]
        codes = hap.reshape(-1)[which[s0:s1, None] * glen + idx]
        err = torch.rand(codes.shape, device=dev, generator=gen) < 0.01
        codes = torch.where(err, (codes + torch.randint(1, 4, codes.shape, device=dev, dtype=torch.uint8, generator=gen)) & 3, codes)
        qq = torch.randint(25, 42, codes.shape, device=dev, dtype=torch.uint8, generator=gen)
        qq = torch.where(err, torch.randint(2, 16, codes.shape, device=dev, dtype=torch.uint8, generator=gen), qq) + 33
        d_b[s0:s1] = acgt[codes.long()]; d_q[s0:s1] = qq
    d_b[::1000, 7] = ord("N"); d_q[::1000, 7] = ord("#")
    off = np.arange(nreads + 1, dtype=np.uint64) * np.uint64(L)
    return d_b.reshape(-1), d_q.reshape(-1), off


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nreads", type=int, default=1_000_000)
    ap.add_argument("--len", type=int, default=100)
    ap.add_argument("--M", type=int, default=2)
    ap.add_argument("--B", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1001)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--glen", type=int, default=0, help="--torchgen: genome length (default: 40x coverage)")
    ap.add_argument("--torchgen", action="store_true", help="generate the reads on the GPU (large collections) and run size-independent checks")
    ap.add_argument("--opt", action="append", default=[], help="ctx option key=value (repeatable)")
    ap.add_argument("--stages", action="store_true", help="print the per-stage CUDA-event times")
    a = ap.parse_args()

    import torch
    from bfqzip_b200 import lib, synth
    dev = torch.device("cuda", 0)
    if a.torchgen:
        d_b, d_q, off = torch_reads(torch, dev, a.nreads, a.len, a.seed, a.glen)
        bases = quals = None
    else:
        bases, quals, off = synth.make_reads(a.nreads, a.len, a.seed)
    nreads = len(off) - 1
    total = int(off[-1]); N = total + nreads
    g = lib.BfqGpu(a.M, a.B, 0)
    if a.stages:
        g.set_option("profile", 1)
    for kv in a.opt:
        k, v = kv.split("=")
        g.set_option(k, int(v))
    P = lib.default_params(eos=0)
    if not a.torchgen:
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
    if a.torchgen:
        # size-independent properties: the eBWT is a permutation of the text, the inverse-LF walk returns every read, and the
        # rebuilt bases differ from the input exactly where noise reduction modified them
        hb = torch.bincount(d_b.long(), minlength=256); hw = torch.bincount(d_bwt[:N].long(), minlength=256)
        hw[P.eos_byte] -= nreads
        nd = int((d_ob[:total] != d_b).sum())
        hq = torch.bincount(d_q.long(), minlength=256); hqs = torch.bincount(d_qs[:N].long(), minlength=256); hqs[P.eos_byte] -= nreads
        lmax = int(d_lcp[:N].max())
        ok = bool(torch.equal(hb, hw)) and nd == st["modified"] and bool(torch.equal(hq, hqs)) and lmax <= a.len
        print(json.dumps({"checks_ok": ok, "bwt_is_permutation": bool(torch.equal(hb, hw)), "qs_is_permutation": bool(torch.equal(hq, hqs)),
                          "bases_changed": nd, "stats_modified": st["modified"], "max_lcp": lmax,
                          "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30}))
        if not ok:
            raise SystemExit(1)
    g.close()


if __name__ == "__main__":
    main()
