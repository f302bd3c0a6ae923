#!/usr/bin/env python
"""Single-collection mode on N GPUs (torchrun, NCCL): every rank checks its eBWT slice and its rebuilt reads against the
ordinary one-GPU pipeline run on its own device, then the mode is timed (device events, max over ranks).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        scripts/dist_check.py [--nreads 1000000] [--len 100] [--steps 3]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG", "WARN")

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nreads", type=int, default=1_000_000)
    ap.add_argument("--len", type=int, default=100)
    ap.add_argument("--M", type=int, default=2)
    ap.add_argument("--B", type=int, default=0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--nocheck", action="store_true")
    a = ap.parse_args()
    import torch
    import torch.distributed as td
    from bfqzip_b200 import dist, lib, synth
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        td.init_process_group("nccl", device_id=torch.device("cuda", lr))
    reads = synth.make_reads(a.nreads, a.len, 1001)
    bases, quals, off = reads
    P = lib.default_params(eos=0)
    eng = dist.GpuEngine(a.M, a.B, lr)
    dreads = eng.upload(reads)
    sl, (ob, oq), (lo, hi), st = dist.run_single_collection(eng, reads, P, rank, world, dreads=dreads)
    ok = True
    if not a.nocheck:
        g = lib.BfqGpu(a.M, a.B, lr)
        (bwt, qs, lcp), (fb, fq), fst = g.pipeline(bases, quals, off, P)
        g.close()
        r0, n = sl["row_base"], sl["n_local"]
        ok &= np.array_equal(sl["bwt"].cpu().numpy(), bwt[r0:r0 + n])
        ok &= np.array_equal(sl["qs"].cpu().numpy(), qs[r0:r0 + n])
        ok &= np.array_equal(sl["lcp"].cpu().numpy(), lcp[r0:r0 + n])
        s, e = int(off[lo]), int(off[hi])
        ok &= np.array_equal(ob.cpu().numpy(), fb[s:e]) and np.array_equal(oq.cpu().numpy(), fq[s:e])
        ok &= all(st[k] == fst[k] for k in dist.STAT_SUM_KEYS)
    times = []
    for _ in range(a.steps):
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        dist.run_single_collection(eng, reads, P, rank, world, dreads=dreads)
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1), 1.0 if ok else 0.0, float(sl["n_local"])], device="cuda", dtype=torch.float64)
        if world > 1:
            tl = [torch.empty_like(t) for _ in range(world)]
            td.all_gather(tl, t)
        else:
            tl = [t]
        times.append(max(float(x[0]) for x in tl))
    if rank == 0:
        ms = min(times)
        print(json.dumps({"mode": "single_collection", "n_gpus": world, "nreads": a.nreads, "len": a.len, "ms_per_step_best": ms,
                          "ms_all": [round(x, 3) for x in times], "gbases_s": int(off[-1]) / ms / 1e6,
                          "slice_rows": [int(x[2]) for x in tl], "parity_ok_all_ranks": all(float(x[1]) == 1.0 for x in tl),
                          "checked": not a.nocheck}), flush=True)
    eng.close()
    if world > 1:
        td.destroy_process_group()
    if not ok:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
