#!/usr/bin/env python
"""Host<->device copy rates with ALL ranks copying at the same time (pinned memory, one process per GPU under torchrun):
the host-side ceiling of the e2e figure at N GPUs.  Rank 0 prints one JSON line with the per-rank and aggregate GB/s.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29577 scripts/pcie_probe_multi.py
"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from bench import _near_gpu, _restore_affinity

rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = 512 << 20
aff = _near_gpu(torch, lr)
h = torch.empty(n, dtype=torch.uint8).pin_memory().zero_(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory().zero_()
_restore_affinity(aff)
d = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def h2d():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)


def both():
    h2d(); d2h()


def timed(fn, reps=6):
    fn(); torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    e0.record()
    s1.wait_event(e0); s2.wait_event(e0)
    for _ in range(reps):
        fn()
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record(); torch.cuda.synchronize()
    return n * reps / e0.elapsed_time(e1) / 1e6          # GB/s per direction


vals = torch.tensor([timed(h2d), timed(d2h), timed(both)], dtype=torch.float64, device="cuda")
if world > 1:
    allv = [torch.empty_like(vals) for _ in range(world)]
    dist.all_gather(allv, vals)
else:
    allv = [vals]
if rank == 0:
    rows = [[round(float(x), 1) for x in v] for v in allv]
    agg = [round(sum(r[i] for r in rows), 1) for i in range(3)]
    print(json.dumps({"n_gpus": world, "bytes_per_copy": n, "per_rank_gbs[h2d,d2h,duplex_each]": rows,
                      "aggregate_gbs": {"h2d_alone": agg[0], "d2h_alone": agg[1], "duplex_each_direction": agg[2]}}), flush=True)
if world > 1:
    dist.destroy_process_group()
