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
    ap.add_argument("--torchgen", action="store_true", help="generate the reads on the GPU (large collections); size-independent checks only")
    ap.add_argument("--opt", action="append", default=[], help="ctx option key=value")
    a = ap.parse_args()
    import torch
    import torch.distributed as td
    from bfqzip_b200 import dist, lib, synth
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    if world > 1:
        td.init_process_group("nccl", device_id=torch.device("cuda", lr))
    P = lib.default_params(eos=0)
    eng = dist.GpuEngine(a.M, a.B, lr, dict(kv.split("=") for kv in a.opt))
    if a.torchgen:
        from prof_step import torch_reads
        d_b, d_q, off = torch_reads(torch, torch.device("cuda", lr), a.nreads, a.len, 1001)
        reads = (None, None, off)
        torch.cuda.empty_cache()          # the library allocates outside torch's cache
        dreads = (d_b, d_q, torch.from_numpy(off.view(np.int64)).to(d_b.device))
    else:
        reads = synth.make_reads(a.nreads, a.len, 1001)
        dreads = eng.upload(reads)
    bases, quals, off = reads
    sl, (ob, oq), (lo, hi), st = dist.run_single_collection(eng, reads, P, rank, world, dreads=dreads)
    ok = True
    props = None
    if a.torchgen and not a.nocheck:
        # size-independent properties, summed over the ranks: the slices tile the row range, the eBWT and QS slices are
        # together a permutation of the text, LCP <= read length, bases differ from the input exactly `modified` times,
        # qualities change only to the smoothing value (mode M=2; the other modes compute the new quality per cluster)
        d_b, d_q, d_off = dreads
        N = int(off[-1]) + len(off) - 1
        s, e = int(off[lo]), int(off[hi])
        hb = torch.bincount(sl["bwt"].long(), minlength=256); hq = torch.bincount(sl["qs"].long(), minlength=256)
        nd = (ob != d_b[s:e]).sum(); ch = oq != d_q[s:e]
        badq = (ch & (oq != ord(">"))).sum()
        v = torch.cat([hb, hq, torch.stack([nd, badq, sl["lcp"].max().long() if sl["n_local"] else nd * 0,
                                            torch.tensor(sl["n_local"], device=nd.device)])]).to(torch.int64)
        rows = torch.tensor([sl["row_base"], sl["n_local"]], device=v.device, dtype=torch.int64)
        if world > 1:
            mx = v[514:515].clone(); td.all_reduce(mx, op=td.ReduceOp.MAX)
            td.all_reduce(v)
            rl = [torch.empty_like(rows) for _ in range(world)]; td.all_gather(rl, rows)
        else:
            mx = v[514:515]; rl = [rows]
        want_b = torch.bincount(d_b.long(), minlength=256); want_b[0] += len(off) - 1
        want_q = torch.bincount(d_q.long(), minlength=256); want_q[0] += len(off) - 1
        spans = sorted((int(x[0]), int(x[1])) for x in rl)
        tiles = spans[0][0] == 0 and all(a0 + n0 == b0 for (a0, n0), (b0, _) in zip(spans, spans[1:])) and spans[-1][0] + spans[-1][1] == N
        props = {"rows_tile": tiles, "bwt_is_permutation": bool(torch.equal(v[:256], want_b)), "qs_is_permutation": bool(torch.equal(v[256:512], want_q)),
                 "bases_changed": int(v[512]), "stats_modified": st["modified"], "bad_quality_changes": int(v[513]), "max_lcp": int(mx[0]), "rows": N}
        ok = tiles and props["bwt_is_permutation"] and props["qs_is_permutation"] and props["bases_changed"] == st["modified"] \
            and (props["bad_quality_changes"] == 0 or a.M != 2) and props["max_lcp"] <= a.len
    elif not a.nocheck:
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
                          "slice_rows": [int(x[2]) for x in tl],
                          # --torchgen: size-independent property checks only; otherwise bit-exact against the one-GPU pipeline
                          ("checks_ok_all_ranks" if a.torchgen else "parity_ok_all_ranks"): all(float(x[1]) == 1.0 for x in tl),
                          "checked": not a.nocheck, "props": props, "stats": st,
                          "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2**30, 1)}), flush=True)
    eng.close()
    if world > 1:
        td.destroy_process_group()
    if not ok:
        raise SystemExit(1)


if __name__ == "__main__":
    main()
