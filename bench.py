#!/usr/bin/env python
"""bench.py -- Gbases/s of eBWT + LCP + smooth + rebuild (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A step is one pass of the whole hot path (parts A+B+C, fused on the device) over one read
collection.  N=1 workload = BASELINE config #2: synthetic 1M reads x 100 bp, M=2, B=0.
N>1: one process per GPU (torchrun), each rank its own block of the same size (block-partitioned
like BFQzip_parallel.py; no data-path collective) -> weak scaling.

`value`  : device-resident inputs/outputs (bfq_pipeline_dev), CUDA events on the launching stream.
`e2e`    : the same metric through the host-buffer C ABI call (bfq_pipeline) with pinned host
           buffers; H2D of the reads and D2H of .bwt/.bwt.qs/.1.lcp + rebuilt reads inside the timing.
`roofline`: the kernel class with the largest share of the timed steps (CUDA events around every launch
           of every class); achieved = that class's algorithmic bytes / its time.
           `roofline_all` lists the same figure for every class (the sort pass and the scans included).
`cpu_baseline` / --impl reference: the UNMODIFIED reference (oracle/_ref: gsufsort + bfq_ext) on the
           host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "Gbases/s eBWT+LCP+smooth+rebuild"
WORKLOAD = dict(nreads=1_000_000, read_len=100, M=2, B=0, seed=1001)


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index
        self.stop_evt = threading.Event()
        self.rows = []

    def _run_nvml(self):
        """NVML from this process: a sample every 2 ms, so that even a 60 ms timed region is covered by dozens."""
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(self.gpu).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID((uuid if uuid.startswith("GPU-") else "GPU-" + uuid).encode())
        mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        bits = ((0x8, 2), (0x40, 3), (0x20, 4), (0x4, 5))      # hw_slowdown, hw_thermal, sw_thermal, sw_power_cap -> row column
        get = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_evt.is_set():
            sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
            r = int(get(h))
            row = [str(sm), str(mx), "", "", "", ""]
            for bit, col in bits:
                row[col] = "Active" if r & bit else "Not Active"
            self.rows.append(row)
            self.stop_evt.wait(0.002)

    def run(self):
        try:
            self._run_nvml()
            return
        except Exception as e:
            print(f"bench.py: NVML clock sampling unavailable ({type(e).__name__}: {e}); falling back to nvidia-smi", file=sys.stderr)
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_evt.wait(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


# ------------------------------------------------------------------------------- reference arm
def run_reference_sample(nreads, read_len, seed, M, B, threads):
    """gsufsort + bfq_ext (unmodified, oracle/_ref) on a sample of the workload.  Returns seconds."""
    from bfqzip_b200 import fastq, synth
    ref = os.path.join(ROOT, "oracle", "_ref")
    bases, quals, off = synth.make_reads(nreads, read_len, seed)
    wd = tempfile.mkdtemp(prefix="bfqbench_")
    try:
        fq = os.path.join(wd, "in.fastq")
        fastq.write_fastq(fq, bases, quals, off)
        t0 = time.perf_counter()
        subprocess.check_call([os.path.join(ref, "gsufsort"), fq, "--bwt", "--qs", "--lcp", "1", "-o", os.path.join(wd, "x")],
                              cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        subprocess.check_call([os.path.join(ref, f"bfq_ext_M{M}_B{B}"), "-e", "x.bwt", "-q", "x.bwt.qs", "-a", "x.1.lcp",
                               "-o", "x", "-l", "250", "-s", "35", "-m", "5", "-T", str(threads)],
                              cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        dt = time.perf_counter() - t0
        assert os.path.getsize(os.path.join(wd, "x.fq")) > 0
        return dt
    finally:
        import shutil
        shutil.rmtree(wd, ignore_errors=True)


def have_ref():
    ref = os.path.join(ROOT, "oracle", "_ref")
    return os.path.exists(os.path.join(ref, "gsufsort")) and os.path.exists(os.path.join(ref, "bfq_ext_M2_B0"))


def run_port_sample(nreads, read_len, seed, M, B):
    """Fallback CPU baseline: the oracle's C restatement (kind 'port'), 1 thread."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from bfqzip_b200 import synth
    bases, quals, off = synth.make_reads(nreads, read_len, seed)
    t0 = time.perf_counter()
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B))
    return time.perf_counter() - t0


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    W = WORKLOAD
    cores = os.cpu_count() or 1
    sample_reads = 60_000
    kind = "reference" if have_ref() else "port"
    times = []
    for it in range(args.warmup + args.steps):
        if kind == "reference":
            dt = run_reference_sample(sample_reads, W["read_len"], W["seed"], W["M"], W["B"], cores)
        else:
            dt = run_port_sample(sample_reads, W["read_len"], W["seed"], W["M"], W["B"])
        if it >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    val = sample_reads * W["read_len"] / t / 1e9
    sample = (f"{sample_reads} reads x {W['read_len']} bp of the {W['nreads']}-read workload (same generator, seed {W['seed']}); "
              f"gsufsort (1 thread, inherent) + bfq_ext -T {cores}" if kind == "reference" else
              f"{sample_reads} reads x {W['read_len']} bp; oracle C port, 1 thread")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "Gbases/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"synthetic {W['nreads']} reads x {W['read_len']} bp single-end, M={W['M']} B={W['B']} (BASELINE config 2)",
                       "sample": sample},
            "cpu_baseline": {"value": val, "unit": "Gbases/s", "cores": cores if kind == "reference" else 1, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)


# ------------------------------------------------------------------------------- GPU arm
def _near_gpu(torch, index):
    """Pins the calling thread to the CPUs NVML lists as closest to CUDA device `index`; returns the previous mask (or None)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(index).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode() if not uuid.startswith("GPU-") else uuid.encode())
        before = os.sched_getaffinity(0)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return before
    except Exception as e:      # affinity is an optimisation only
        print(f"bench.py: no GPU-local CPU affinity ({type(e).__name__}: {e})", file=sys.stderr)
        return None


def _restore_affinity(before):
    if before:
        try:
            os.sched_setaffinity(0, before)
        except OSError:
            pass


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from bfqzip_b200 import lib, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this benchmark has no CPU fallback for the product arm")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ["NCCL_DEBUG"] = "WARN"        # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W = WORKLOAD
    M, B = W["M"], W["B"]
    bases, quals, off = synth.make_reads(W["nreads"], W["read_len"], W["seed"] + rank)
    nreads = len(off) - 1
    total = int(off[-1]); N = total + nreads
    P = lib.default_params(eos=0)           # the driver's spelling: -s 0 -m 5 -l 250

    g = lib.BfqGpu(M, B, local)
    stream = torch.cuda.current_stream()
    g.set_stream(stream.cuda_stream)
    g.set_option("profile", 1)

    dev = torch.device("cuda", local)
    d_bases = torch.from_numpy(bases).to(dev); d_quals = torch.from_numpy(quals).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_bwt = torch.empty(N + 64, dtype=torch.uint8, device=dev); d_qs = torch.empty_like(d_bwt); d_lcp = torch.empty_like(d_bwt)
    d_ob = torch.empty(total + 64, dtype=torch.uint8, device=dev); d_oq = torch.empty_like(d_ob)

    def step_dev():
        return g.pipeline_dev(d_bases, d_quals, d_off, nreads, total, P, d_bwt, d_qs, d_lcp, d_ob, d_oq)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        st = step_dev()
    barrier()
    g.prof_reset()
    l0 = g.launch_count()
    sampler = ClockSampler(local); sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        st = step_dev()
    e1.record(stream)
    barrier()
    sampler.stop_evt.set(); sampler.join()
    ms = e0.elapsed_time(e1)
    launches = g.launch_count() - l0
    prof = {k: g.prof_get(k) for k in ("radix_pass", "radix_hist", "pack", "keys", "group", "defer", "hist", "rank",
                                        "cluster", "cluster_long", "rule11", "walk", "assemble", "unpermute")}

    # ---- e2e: host pinned buffers through bfq_pipeline (H2D + D2H inside the timing)
    g.set_option("profile", 0)
    # the pinned buffers are first touched from the cores next to this rank's GPU, so they sit on its NUMA node
    # (several ranks copying through one socket halve each other's PCIe rate); the affinity is restored afterwards
    numa = _near_gpu(torch, local)
    h_b = torch.from_numpy(bases).pin_memory(); h_q = torch.from_numpy(quals).pin_memory()
    h_off = torch.from_numpy(off.view(np.int64)).pin_memory()
    h_bwt = torch.empty(N, dtype=torch.uint8).pin_memory(); h_qs = torch.empty(N, dtype=torch.uint8).pin_memory()
    h_lcp = torch.empty(N, dtype=torch.uint8).pin_memory()
    h_ob = torch.empty(total, dtype=torch.uint8).pin_memory(); h_oq = torch.empty(total, dtype=torch.uint8).pin_memory()

    h_bwt.zero_(); h_qs.zero_(); h_lcp.zero_(); h_ob.zero_(); h_oq.zero_()
    _restore_affinity(numa)

    def step_host():
        return g.pipeline(h_b, h_q, h_off, P, out=(h_bwt, h_qs, h_lcp, h_ob.numpy(), h_oq.numpy()))

    for _ in range(max(1, min(args.warmup, 2))):
        step_host()
    barrier()
    t0 = time.perf_counter()
    f0 = torch.cuda.Event(enable_timing=True); f1 = torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    for _ in range(args.steps):
        step_host()
    f1.record(stream)
    barrier()
    ms_e2e = f0.elapsed_time(f1)            # device clock; each call is synchronous on this stream
    wall_e2e = (time.perf_counter() - t0) * 1e3
    # sanity: the e2e result equals the device-resident result
    assert torch.equal(h_ob.to(dev), d_ob[:total]) and torch.equal(h_oq.to(dev), d_oq[:total])

    # ---- max over ranks
    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    bases_all = total * world * args.steps
    value = bases_all / (ms * 1e-3) / 1e9
    e2e_val = bases_all / (ms_e2e * 1e-3) / 1e9

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        KNAME = {"radix_pass": "rs_onesweep_kernel", "radix_hist": "rs_hist_kernel", "pack": "pack_reads_kernel+pack3_kernel",
                 "keys": "rs_suffix_keys_kernel", "group": "group_tile_kernel", "defer": "defer_*_kernel", "hist": "hist256_kernel",
                 "rank": "rank_*_kernel", "cluster": "cluster_tile_kernel", "cluster_long": "cluster_long_kernel",
                 "rule11": "rule11_kernel", "walk": "walk_kernel", "assemble": "assemble_kernel / compact_reads_kernel",
                 "unpermute": "rs_onesweep_kernel<u32,u16> + rs_place_kernel"}
        # ncu dram__bytes_read.sum + dram__bytes_write.sum per launch (profiles/r1k_hot_kernels.txt: the device-resident step,
        # flat mode -- the group kernel also writes the suffix array, the cluster kernel one stored-quality byte per row)
        TRAFFIC = {"group": 1948.5e6, "radix_pass": 2409.5e6, "cluster": 417.8e6, "keys": 1393.9e6}
        roof_all = {}
        for k, (k_ms, k_n, k_bytes) in prof.items():
            if k_ms > 0 and k_n:
                roof_all[k] = {"kernel": KNAME[k], "ms_per_step": k_ms / args.steps, "launches_per_step": k_n / args.steps,
                               "achieved": (k_bytes / 1e9) / (k_ms * 1e-3), "frac": (k_bytes / 1e9) / (k_ms * 1e-3) / peak,
                               "share_of_step": k_ms / ms}
        dom = max(roof_all, key=lambda k: roof_all[k]["ms_per_step"])
        rp_ms, rp_n, rp_bytes = prof[dom]
        achieved = roof_all[dom]["achieved"]
        # CPU baseline on a bounded sample (rank 0, N=1 only)
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            sample_reads = 40_000
            if have_ref():
                dt = run_reference_sample(sample_reads, W["read_len"], W["seed"], M, B, 1)
                kind, cores, how = "reference", 1, "gsufsort + bfq_ext -T 1 (what BFQzip_ext.py runs), unmodified reference binaries"
            else:
                dt = run_port_sample(sample_reads, W["read_len"], W["seed"], M, B)
                kind, cores, how = "port", 1, "oracle C port"
            cpu = {"value": sample_reads * W["read_len"] / dt / 1e9, "unit": "Gbases/s", "cores": cores, "kind": kind,
                   "sample": f"{sample_reads} reads x {W['read_len']} bp of the workload (seed {W['seed']}), {dt:.1f} s; {how}"}
        line = {
            "metric": METRIC, "value": value, "unit": "Gbases/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"synthetic {W['nreads']} reads x {W['read_len']} bp single-end per GPU, M={M} B={B}, K=16 m=5 eos=0 "
                                   f"(BASELINE config 2; seeded generator bfqzip_b200/synth.py, seed {W['seed']}+rank)",
                       "rows_per_gpu": N, "parallelism": f"block-per-gpu x{world}, no collective",
                       "l2": "inputs (200 MB) and every intermediate array exceed the 126 MB L2; no explicit flush"},
            "e2e": {"value": e2e_val, "unit": "Gbases/s", "h2d_bytes_per_step": int(2 * total + 8 * (nreads + 1)),
                    "d2h_bytes_per_step": int(3 * N + 2 * total), "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": f"{KNAME[dom]} ({dom})", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak if peak else None, "traffic": TRAFFIC.get(dom),
                         "peak_source": peak_src, "launches": rp_n, "avg_launch_ms": rp_ms / rp_n if rp_n else None,
                         "share_of_step": rp_ms / ms if ms else None},
            "roofline_all": roof_all,
            "stages_ms_per_step": {k: v[0] / args.steps for k, v in prof.items()},
            "cpu_baseline": cpu,
            "clocks": sampler.summary(),
            "stats": st,
        }
        _emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: while the arms run, file descriptor 1 points at stderr (libraries such as
    # NCCL write banners to it from native code); print() is given the real stdout back through _emit
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    global _OUT
    _OUT = os.fdopen(real, "w")
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


_OUT = None


def _emit(line):
    out = _OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


if __name__ == "__main__":
    main()
