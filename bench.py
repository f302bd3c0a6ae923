#!/usr/bin/env python
"""bench.py -- Gbases/s of eBWT + LCP + smooth + rebuild (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A step is one pass of the whole hot path (parts A+B+C, fused on the device) over the workload.

N = 1   headline = BASELINE config 2 (synthetic 1M reads x 100 bp, M=2 B=0) on one B200.  Secondary records in the same line:
        `config4_blocks` (BASELINE config 4, all 8 blocks one after the other on the one GPU: the base of the strong-scaling
        series) and `single_collection` (BASELINE config 5 at 1/8 size through the single-collection code path, one rank).
N > 1   headline = BASELINE config 4: paired 2 x 20M x 150 bp, M=3, EIGHT FIXED BLOCKS cut as split_fastq / split_fastq_2 cut them
        (BFQzip_parallel.py:288-360) and dealt to the ranks in contiguous runs (8/4/2/1 blocks per GPU), no data-path
        collective -> "scaling": "strong", value = 6e9 bases / slowest rank.  Secondary: `weak_config2` (one config-2
        collection per GPU: round 1's headline) and `single_collection` (BASELINE config 5 scaled to N/8 of its 50M reads:
        ONE eBWT over all ranks, one NCCL all-to-all, with its stage split).

`value`  : device-resident inputs/outputs (bfq_pipeline_dev), CUDA events on the launching stream, max over ranks.
`e2e`    : the same workload through the host-buffer C ABI call (bfq_pipeline) with pinned host buffers; H2D of the reads and
           D2H of .bwt/.bwt.qs/.1.lcp + rebuilt reads inside the timing.
`roofline`: the kernel class with the largest share of the timed steps (CUDA events around every launch of every class);
           achieved = that class's DRAM-level algorithmic bytes / its time; `roofline_all` lists every class.
`cpu_baseline` / --impl reference: the UNMODIFIED reference binaries (oracle/_ref) on the host cores.
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
CFG2 = dict(nreads=1_000_000, read_len=100, M=2, B=0, seed=1001)
CFG4 = dict(npairs=20_000_000, read_len=150, M=3, B=0, seed=1003, coverage=40.0, fragment=350, blocks=8)
CFG5 = dict(nreads=50_000_000, read_len=150, M=0, B=0, seed=1004, coverage=40.0)
WORKLOAD = CFG2
PROF_KEYS = ("radix_pass", "radix_hist", "pack", "keys", "group", "defer", "hist", "rank", "cluster", "cluster_long", "rule11",
             "walk", "assemble", "unpermute")
KNAME = {"radix_pass": "rs_onesweep_kernel", "radix_hist": "rs_hist_kernel", "pack": "pack_reads_kernel+pack3_kernel",
         "keys": "rs_suffix_keys_kernel / rs_select_kernel", "group": "group_tile_kernel", "defer": "defer_*_kernel", "hist": "hist256_kernel",
         "rank": "rank_*_kernel", "cluster": "cluster_tile_kernel", "cluster_long": "cluster_long_kernel",
         "rule11": "rule11_kernel", "walk": "walk_kernel / emit_*_pairs_kernel", "assemble": "assemble_kernel / compact_reads_kernel",
         "unpermute": "rs_onesweep_kernel<u32,u16> + rs_place_kernel"}
# what bounds each class: "hbm" = streaming, the fraction of the HBM peak is the figure of merit;
# "issue" = instruction issue + L2 gathers, the fraction of the HBM peak is reported but is not what limits the kernel
BOUND = {"group": "issue", "walk": "dram-sector-rate", "rule11": "dram-sector-rate"}


def workload_cfg2():
    W = CFG2
    return (f"BASELINE config 2: synthetic {W['nreads']} reads x {W['read_len']} bp single-end, M={W['M']} B={W['B']}, K=16 m=5 "
            f"(seeded generator bfqzip_b200/synth.py make_reads, seed {W['seed']})")


def workload_cfg4(npairs=None):
    V = CFG4
    npairs = npairs or V["npairs"]
    return (f"BASELINE config 4: synthetic paired 2 x {npairs} reads x {V['read_len']} bp, M={V['M']} B={V['B']}, K=16 m=5, {V['blocks']} fixed blocks "
            f"cut as split_fastq / split_fastq_2 do (counter-based generator bfqzip_b200/synth.py hash_paired, seed {V['seed']}, coverage {V['coverage']:g}, fragment {V['fragment']})")


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




# ------------------------------------------------------------------------------- reference arm (CPU, oracle/_ref)
REF = os.path.join(ROOT, "oracle", "_ref")


def have_ref():
    return os.path.exists(os.path.join(REF, "gsufsort")) and os.path.exists(os.path.join(REF, "bfq_ext_M2_B0"))


def _write_fixed(path, bases, quals, nreads, read_len):
    from bfqzip_b200 import fastq
    with open(path, "wb") as f:
        f.write(fastq.format_fastq_fixed(bases, quals, nreads, read_len))


def ref_pipeline(fq, wd, M, B, threads, step1="gsufsort"):
    """The reference's two tools on one FASTQ in its own directory: step 1 (gsufsort, or the eGap driver as BFQzip_ext.py:173-177
    spells it) + bfq_ext as BFQzip_ext.py:207-215 spells it.  Returns (seconds step 1, seconds step 3)."""
    t0 = time.perf_counter()
    if step1 == "egap":
        mb = max(1, os.path.getsize(fq) >> 20)
        subprocess.check_call([sys.executable, os.path.join(REF, "egap", "eGap"), fq, "--em", "--mem", str(mb), "--qs", "-o",
                               os.path.join(wd, "x"), "--lcp", "--lbytes", "1"], cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        eos = "0"
    else:
        subprocess.check_call([os.path.join(REF, "gsufsort"), fq, "--bwt", "--qs", "--lcp", "1", "-o", os.path.join(wd, "x")],
                              cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        eos = "35"
    t1 = time.perf_counter()
    subprocess.check_call([os.path.join(REF, f"bfq_ext_M{M}_B{B}"), "-e", "x.bwt", "-q", "x.bwt.qs", "-a", "x.1.lcp",
                           "-o", "x", "-l", "250", "-s", eos, "-m", "5", "-T", str(threads)],
                          cwd=wd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t2 = time.perf_counter()
    assert os.path.getsize(os.path.join(wd, "x.fq")) > 0
    return t1 - t0, t2 - t1


def ref_cfg2(nreads, threads, step1="gsufsort", blocks=1):
    """Config 2's first `nreads` reads through the reference (`blocks` > 1: that many concurrent block pipelines, the
    BFQzip_parallel.py -t n semantics, each with threads // blocks bfq_ext threads).  Returns wall seconds (tools only)."""
    import shutil
    from concurrent.futures import ThreadPoolExecutor
    from bfqzip_b200 import parallel, synth
    W = CFG2
    bases, quals, off = synth.make_reads(W["nreads"], W["read_len"], W["seed"])
    L = W["read_len"]
    root = tempfile.mkdtemp(prefix="bfqbench_")
    try:
        jobs = []
        for i, (s, e) in enumerate(parallel.split_blocks(nreads, blocks)):
            wd = os.path.join(root, f"b{i}"); os.makedirs(wd)
            fq = os.path.join(wd, "in.fastq")
            _write_fixed(fq, bases[s * L:e * L], quals[s * L:e * L], e - s, L)
            jobs.append((fq, wd))
        per = max(1, threads // len(jobs))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(len(jobs)) as ex:
            parts = list(ex.map(lambda j: ref_pipeline(j[0], j[1], W["M"], W["B"], per, step1), jobs))
        return time.perf_counter() - t0, parts
    finally:
        shutil.rmtree(root, ignore_errors=True)


def ref_cfg4_sample(pairs_per_block, threads):
    """Config 4's eight blocks, each cut down to its first `pairs_per_block` pairs, as eight concurrent block pipelines."""
    import shutil
    from concurrent.futures import ThreadPoolExecutor
    from bfqzip_b200 import parallel, synth
    W = CFG4
    g = synth.hash_genome_len(2 * W["npairs"], W["read_len"], W["coverage"], W["fragment"])
    L = W["read_len"]
    root = tempfile.mkdtemp(prefix="bfqbench_")
    try:
        jobs = []
        for i, (lo, hi) in enumerate(parallel.split_blocks(W["npairs"], W["blocks"])):
            hi = lo + pairs_per_block
            m = [synth.hash_reads(W["npairs"], L, W["seed"], g, mate, W["fragment"], lo, hi) for mate in (0, 1)]
            wd = os.path.join(root, f"b{i}"); os.makedirs(wd)
            fq = os.path.join(wd, "in.fastq")
            _write_fixed(fq, np.concatenate([m[0][0], m[1][0]]), np.concatenate([m[0][1], m[1][1]]), 2 * (hi - lo), L)
            jobs.append((fq, wd))
        per = max(1, threads // len(jobs))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(len(jobs)) as ex:
            list(ex.map(lambda j: ref_pipeline(j[0], j[1], W["M"], W["B"], per), jobs))
        return time.perf_counter() - t0
    finally:
        shutil.rmtree(root, ignore_errors=True)


def run_port_sample(nreads, read_len, seed, M, B):
    """Fallback CPU baseline when oracle/_ref is absent: the oracle's C restatement (kind 'port'), 1 thread."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H
    from bfqzip_b200 import synth
    bases, quals, off = synth.make_reads(nreads, read_len, seed)
    t0 = time.perf_counter()
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 35)
    H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B))
    return time.perf_counter() - t0


def host_info():
    model = ""
    try:
        with open("/proc/cpuinfo") as f:
            model = next((l.split(":", 1)[1].strip() for l in f if l.startswith("model name")), "")
        with open("/proc/meminfo") as f:
            kb = int(next(l for l in f if l.startswith("MemTotal")).split()[1])
    except Exception:
        kb = 0
    return {"nproc": os.cpu_count() or 1, "cpu": model, "ram_gb": round(kb / 2**20, 1)}


def reference_arm(args):
    """The reference's own CPU implementation of the path on this box's host cores.  Each configuration is a deterministic
    CPU run of a minute or more, so it is timed ONCE and that time is reported for every one of the K steps."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    quick = int(os.environ.get("BFQ_BENCH_QUICK", "0"))          # development only: shrink the runs
    W = CFG2
    variants = {}
    if args.gpus <= 1:
        workload = workload_cfg2()
        if have_ref():
            kind, used = "reference", cores
            n_full = W["nreads"] // (quick or 1)
            t, parts = ref_cfg2(n_full, cores)
            bases_done = n_full * W["read_len"]
            sample = (f"{'full config 2' if not quick else 'QUICK sample'}: {n_full} reads x {W['read_len']} bp (seed {W['seed']}), timed once: gsufsort "
                      f"{parts[0][0]:.1f} s (1 thread, inherent) + bfq_ext -T {cores} {parts[0][1]:.1f} s; unmodified reference binaries (oracle/_ref)")
            # the other two legs SURVEY 8(d) lists: what the driver really runs (eGap + 1 thread) on a stated sample -- eGap alone
            # needs ~10 minutes for the full collection -- and the whole collection as eight concurrent block pipelines
            # BFQ_BENCH_EGAP_FULL=1: the whole collection through eGap too (minutes; run once per pool, recorded in profiles/)
            n_eg = (W["nreads"] if os.environ.get("BFQ_BENCH_EGAP_FULL") else 40_000) // (quick or 1)
            te, pe = ref_cfg2(n_eg, 1, step1="egap")
            variants["egap_bfq_ext_T1"] = {"value": n_eg * W["read_len"] / te / 1e9, "unit": "Gbases/s", "cores": 1,
                                           "sample": f"first {n_eg} reads of config 2 ({100.0 * n_eg / W['nreads']:.0f} %): eGap --em {pe[0][0]:.1f} s + bfq_ext -T 1 {pe[0][1]:.1f} s (BFQzip_ext.py:173-177, 207-215)"}
            tb, _ = ref_cfg2(n_full, cores, blocks=8)
            variants["blocks8_concurrent"] = {"value": bases_done / tb / 1e9, "unit": "Gbases/s", "cores": cores,
                                              "sample": f"{n_full} reads as 8 concurrent block pipelines (BFQzip_parallel.py:102-119 semantics), gsufsort + bfq_ext -T {max(1, cores // 8)} each, {tb:.1f} s"}
        else:
            kind, used = "port", 1
            n_full = 60_000
            t = run_port_sample(n_full, W["read_len"], W["seed"], W["M"], W["B"])
            bases_done = n_full * W["read_len"]
            sample = f"{n_full} reads x {W['read_len']} bp of config 2; oracle C port, 1 thread (oracle/_ref not built on this box)"
        scaling = "weak"
    else:
        V = CFG4
        workload = workload_cfg4()
        if not have_ref():
            _emit({"impl": "reference", "unavailable": "oracle/_ref is not built on this box; config 4 has no port leg"})
            return
        kind, used = "reference", cores
        ppb = 25_000 // (quick or 1)
        t = ref_cfg4_sample(ppb, cores)
        bases_done = V["blocks"] * 2 * ppb * V["read_len"]
        sample = (f"1/{V['npairs'] // V['blocks'] // ppb} of every block: {V['blocks']} concurrent block pipelines of 2 x {ppb} reads x {V['read_len']} bp "
                  f"(gsufsort + bfq_ext -T {max(1, cores // 8)} each), timed once; a full block needs ~15 minutes of CPU")
        scaling = "strong"
    val = bases_done / t / 1e9
    cpu = {"value": val, "unit": "Gbases/s", "cores": used, "kind": kind, "sample": sample, "host": host_info()}
    if variants:
        cpu["variants"] = variants
    _emit({"impl": "reference", "metric": METRIC, "value": val, "unit": "Gbases/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "timed_runs": 1, "higher_is_better": True,
           "scaling": scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
           # N = 1: the whole configuration is timed, the config equals the product arm's; N > 1: a stated sample of it
           "config": ({"workload": workload} if args.gpus <= 1 and kind == "reference" and not quick else {"workload": workload, "sample": sample}),
           "cpu_baseline": cpu,
           "e2e": {"value": val, "unit": "Gbases/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0})


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


class Env:
    """torch / torch.distributed / rank bookkeeping shared by the three measurements."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; this benchmark has no CPU fallback for the product arm")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.current_stream()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def sum_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t)
        return [float(x) for x in t]

    def timed(self, fn, steps):
        """barrier + synchronize, K calls between two events on the launching stream, barrier + synchronize.  Returns ms."""
        torch = self.torch
        self.barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for _ in range(steps):
            fn()
        e1.record(self.stream)
        self.barrier()
        return e0.elapsed_time(e1)


def roofline_table(prof, steps, ms_total, peak):
    out = {}
    for k, (k_ms, k_n, k_bytes) in prof.items():
        if k_ms > 0 and k_n:
            gbs = (k_bytes / 1e9) / (k_ms * 1e-3)
            out[k] = {"kernel": KNAME[k], "bound": BOUND.get(k, "hbm"), "ms_per_step": k_ms / steps, "launches_per_step": k_n / steps,
                      "achieved": gbs, "frac": gbs / peak, "share_of_step": k_ms / ms_total}
    return out


def bench_cfg2(env, steps, warmup):
    """One config-2 collection per rank (device-resident + host-buffer call)."""
    torch = env.torch
    from bfqzip_b200 import lib, synth
    W = CFG2
    M, B = W["M"], W["B"]
    bases, quals, off = synth.make_reads(W["nreads"], W["read_len"], W["seed"] + env.rank)
    nreads = len(off) - 1
    total = int(off[-1]); N = total + nreads
    P = lib.default_params(eos=0)           # the driver's spelling: -s 0 -m 5 -l 250
    g = lib.BfqGpu(M, B, env.local)
    g.set_stream(env.stream.cuda_stream)
    g.set_option("profile", 1)
    dev = env.dev
    d_bases = torch.from_numpy(bases).to(dev); d_quals = torch.from_numpy(quals).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_bwt = torch.empty(N + 64, dtype=torch.uint8, device=dev); d_qs = torch.empty_like(d_bwt); d_lcp = torch.empty_like(d_bwt)
    d_ob = torch.empty(total + 64, dtype=torch.uint8, device=dev); d_oq = torch.empty_like(d_ob)
    res = {}

    def step_dev():
        res["st"] = g.pipeline_dev(d_bases, d_quals, d_off, nreads, total, P, d_bwt, d_qs, d_lcp, d_ob, d_oq)

    for _ in range(warmup):
        step_dev()
    env.barrier()
    g.prof_reset()
    l0 = g.launch_count()
    sampler = ClockSampler(env.local); sampler.start()
    ms = env.timed(step_dev, steps)
    sampler.stop_evt.set(); sampler.join()
    launches = g.launch_count() - l0
    prof = {k: g.prof_get(k) for k in PROF_KEYS}

    # ---- e2e: host pinned buffers through bfq_pipeline (H2D + D2H inside the timing)
    g.set_option("profile", 0)
    # the pinned buffers are first touched from the cores next to this rank's GPU, so they sit on its NUMA node
    numa = _near_gpu(torch, env.local)
    h_b = torch.from_numpy(bases).pin_memory(); h_q = torch.from_numpy(quals).pin_memory()
    h_off = torch.from_numpy(off.view(np.int64)).pin_memory()
    pin = lambda k: torch.empty(k, dtype=torch.uint8).pin_memory().zero_()
    h_bwt, h_qs, h_lcp, h_ob, h_oq = pin(N), pin(N), pin(N), pin(total), pin(total)
    _restore_affinity(numa)

    def step_host():
        g.pipeline(h_b, h_q, h_off, P, out=(h_bwt, h_qs, h_lcp, h_ob.numpy(), h_oq.numpy()))

    for _ in range(max(1, min(warmup, 2))):
        step_host()
    ms_e2e = env.timed(step_host, steps)        # device clock; each call is synchronous on this stream
    same = bool(torch.equal(h_ob.to(dev), d_ob[:total]) and torch.equal(h_oq.to(dev), d_oq[:total]))

    # the same call for a caller who wants only what the reference's steps 1-3 deliver in the end -- the smoothed reads:
    # bwt / qs / lcp = NULL, the three intermediate files (gsufsort's output, bfq_ext's input) never leave the GPU
    def step_host_reads():
        g.pipeline(h_b, h_q, h_off, P, out=(None, None, None, h_ob.numpy(), h_oq.numpy()))

    h_ob.zero_(); h_oq.zero_()
    step_host_reads()
    same_reads = bool(torch.equal(h_ob.to(dev), d_ob[:total]) and torch.equal(h_oq.to(dev), d_oq[:total]))
    ms_e2e_reads = env.timed(step_host_reads, steps)
    g.close()
    ms, ms_e2e, ms_e2e_reads = env.max_over_ranks([ms, ms_e2e, ms_e2e_reads])
    return dict(ms=ms, ms_e2e=ms_e2e, ms_e2e_reads=ms_e2e_reads, d2h_reads=int(2 * total) * env.world, e2e_reads_equal=same_reads, steps=steps, bases_per_step=total * env.world, launches=int(launches), prof=prof, stats=res["st"],
                clocks=sampler.summary(), rows_per_gpu=N, h2d=int(2 * total + 8 * (nreads + 1)) * env.world,
                d2h=int(3 * N + 2 * total) * env.world, e2e_equals_device=same,
                workload=workload_cfg2() + (f"; one such collection per GPU, seed {W['seed']}+rank" if env.world > 1 else ""))


def bench_cfg4(env, steps, warmup):
    """BASELINE config 4: paired collection in EIGHT FIXED BLOCKS (parallel.split_blocks = split_fastq, block i = mates 1 of
    its pairs followed by their mates 2 = split_fastq_2), dealt to the ranks by parallel.blocks_of_rank (8/4/2/1 per GPU).
    Every block is its own collection with its own eBWT; no data-path collective.  The result depends on the block count
    only, so the runs on 1, 2, 4 and 8 GPUs do the same work: strong scaling."""
    torch = env.torch
    from bfqzip_b200 import lib, parallel, synth
    V = CFG4
    quick = int(os.environ.get("BFQ_BENCH_QUICK", "0"))
    npairs = V["npairs"] // (quick or 1)
    L, M, B = V["read_len"], V["M"], V["B"]
    glen = synth.hash_genome_len(2 * npairs, L, V["coverage"], V["fragment"])
    cuts = parallel.split_blocks(npairs, V["blocks"])
    mine = list(parallel.blocks_of_rank(len(cuts), env.rank, env.world))
    P = lib.default_params(eos=0)
    dev = env.dev
    g = lib.BfqGpu(M, B, env.local)
    g.set_stream(env.stream.cuda_stream)
    g.set_option("profile", 1)
    blocks = []
    for b in mine:
        lo, hi = cuts[b]
        d_b, d_q, off = synth.hash_paired_block_torch(torch, dev, npairs, L, V["seed"], glen, V["fragment"], lo, hi)
        nreads = len(off) - 1; total = int(off[-1]); N = total + nreads
        u8 = lambda k: torch.empty(k + 64, dtype=torch.uint8, device=dev)
        blocks.append(dict(b=b, d_b=d_b, d_q=d_q, off=off, d_off=torch.from_numpy(off.view(np.int64)).to(dev), nreads=nreads,
                           total=total, N=N, d_bwt=u8(N), d_qs=u8(N), d_lcp=u8(N), d_ob=u8(total), d_oq=u8(total)))
    torch.cuda.synchronize()
    torch.cuda.empty_cache()                      # the library allocates outside torch's cache

    def step_dev():
        for k in blocks:
            k["st"] = g.pipeline_dev(k["d_b"], k["d_q"], k["d_off"], k["nreads"], k["total"], P, k["d_bwt"], k["d_qs"], k["d_lcp"], k["d_ob"], k["d_oq"])

    for _ in range(warmup):
        step_dev()
    env.barrier()
    g.prof_reset()
    l0 = g.launch_count()
    sampler = ClockSampler(env.local); sampler.start()
    ms = env.timed(step_dev, steps)
    sampler.stop_evt.set(); sampler.join()
    launches = g.launch_count() - l0
    prof = {k: g.prof_get(k) for k in PROF_KEYS}
    # size-independent checks on every block: the rebuilt bases differ from the input exactly `modified` times, the eBWT
    # is a permutation of the block's symbols, LCP <= read length
    ok = True
    for k in blocks:
        nd = int((k["d_ob"][:k["total"]] != k["d_b"]).sum())
        hb = torch.bincount(k["d_b"].long(), minlength=256); hw = torch.bincount(k["d_bwt"][:k["N"]].long(), minlength=256)
        hw[P.eos_byte] -= k["nreads"]
        ok = ok and nd == k["st"]["modified"] and bool(torch.equal(hb, hw)) and int(k["d_lcp"][:k["N"]].max()) <= L and k["st"]["nreads"] == k["nreads"]
    # ---- e2e: every block through the host-buffer call; the inputs of all blocks of this rank sit in pinned host memory,
    # the outputs of a block land in ONE pinned output set (a caller writes it to the block's files before the next block)
    g.set_option("profile", 0)
    numa = _near_gpu(torch, env.local)
    for k in blocks:
        k["h_b"] = k["d_b"].cpu().pin_memory(); k["h_q"] = k["d_q"].cpu().pin_memory()
        k["h_off"] = torch.from_numpy(k["off"].view(np.int64)).pin_memory()
    Nmax = max(k["N"] for k in blocks); Tmax = max(k["total"] for k in blocks)
    pin = lambda n: torch.empty(n, dtype=torch.uint8).pin_memory().zero_()
    h_bwt, h_qs, h_lcp, h_ob, h_oq = pin(Nmax), pin(Nmax), pin(Nmax), pin(Tmax), pin(Tmax)
    _restore_affinity(numa)

    def step_host():
        for k in blocks:
            g.pipeline(k["h_b"], k["h_q"], k["h_off"], P, out=(h_bwt[:k["N"]], h_qs[:k["N"]], h_lcp[:k["N"]], h_ob.numpy(), h_oq.numpy()))

    step_host()
    ms_e2e = env.timed(step_host, steps)
    last = blocks[-1]
    same = bool(torch.equal(h_ob[:last["total"]].to(dev), last["d_ob"][:last["total"]]) and torch.equal(h_bwt[:last["N"]].to(dev), last["d_bwt"][:last["N"]]))
    g.close()
    stats = {}
    for k in blocks:
        for key, v in k["st"].items():
            stats[key] = stats.get(key, 0) + v
    keys = sorted(stats)
    summed = env.sum_over_ranks([float(stats[key]) for key in keys] + [1.0 if ok else 0.0, 1.0 if same else 0.0])
    ms, ms_e2e = env.max_over_ranks([ms, ms_e2e])
    total_all = 2 * npairs * L
    N_all = total_all + 2 * npairs
    return dict(ms=ms, ms_e2e=ms_e2e, steps=steps, bases_per_step=total_all, launches=int(launches), prof=prof,
                stats={key: int(v) for key, v in zip(keys, summed)}, checks_ok=summed[-2] == env.world, e2e_equals_device=summed[-1] == env.world,
                clocks=sampler.summary(), rows_per_gpu=sum(k["N"] for k in blocks), blocks_per_gpu=len(blocks),
                h2d=int(2 * total_all + 8 * (2 * npairs + V["blocks"])), d2h=int(3 * N_all + 2 * total_all),
                workload=workload_cfg4(npairs))


def bench_single(env, steps, warmup):
    """BASELINE config 5 scaled to N/8 of its 50M reads: ONE eBWT over all ranks (bfqzip_b200/dist.py): replicated reads,
    key-range selection, per-rank sort + group ordering + smoothing, ONE NCCL all-to-all of the smoothed rows, placement."""
    torch = env.torch
    from bfqzip_b200 import dist as bdist, lib, synth
    S = CFG5
    quick = int(os.environ.get("BFQ_BENCH_QUICK", "0"))
    nreads = S["nreads"] * env.world // 8 // (quick or 1)
    L, M, B = S["read_len"], S["M"], S["B"]
    glen = synth.hash_genome_len(nreads, L, S["coverage"], L)
    dev = env.dev
    d_b, d_q = synth.hash_reads_torch(torch, dev, nreads, L, S["seed"], glen)        # every rank generates the same collection
    off = np.arange(nreads + 1, dtype=np.uint64) * np.uint64(L)
    dreads = (d_b, d_q, torch.from_numpy(off.view(np.int64)).to(dev))
    reads = (None, None, off)
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    eng = bdist.GpuEngine(M, B, env.local)
    eng.g.set_stream(env.stream.cuda_stream)
    eng.g.set_option("profile", 1)
    P = lib.default_params(eos=0)
    res, timing = {}, {}

    def step():
        res["out"] = bdist.run_single_collection(eng, reads, P, env.rank, env.world, dreads=dreads, timing=timing)

    for _ in range(warmup):
        step()
    env.barrier()
    eng.g.prof_reset(); timing.clear()
    l0 = eng.g.launch_count()
    ms = env.timed(step, steps)
    launches = eng.g.launch_count() - l0
    prof = {k: eng.g.prof_get(k) for k in PROF_KEYS}
    sl, (ob, oq), (lo, hi), st = res["out"]
    # size-independent checks, summed over the ranks: the slices tile the row range, the rebuilt bases of this rank's reads
    # differ from the input exactly as often as the slices report modifications (summed), LCP <= read length
    s, e = int(off[lo]), int(off[hi])
    nd = int((ob != d_b[s:e]).sum())
    lmax = int(sl["lcp"].max()) if sl["n_local"] else 0
    tot = env.sum_over_ranks([float(nd), float(sl["n_local"])])
    mx = env.max_over_ranks([float(lmax), ms] + [timing.get(k, 0.0) * 1e3 / steps for k in ("build", "exchange", "place")] +
                            [prof[k][0] / steps for k in PROF_KEYS])
    N = int(off[-1]) + nreads
    ok = int(tot[0]) == st["modified"] and int(tot[1]) == N and mx[0] <= L
    eng.close()
    return dict(ms=mx[1], steps=steps, bases_per_step=int(off[-1]), launches=int(launches), stats=st, checks_ok=bool(ok), rows=N,
                wide_positions=N >= (1 << 32) - 64,
                stage_ms_per_step={"build (select + sort + group + smooth + emit)": mx[2], "exchange (all-to-all)": mx[3], "place": mx[4]},
                kernel_class_ms_per_step={k: round(v, 3) for k, v in zip(PROF_KEYS, mx[5:]) if v > 0},
                workload=(f"synthetic {nreads} reads x {L} bp, ONE collection over {env.world} GPU(s), M={M} B={B} (BASELINE config 5 scaled to "
                          f"{env.world}/8 of its {S['nreads']} reads; counter-based generator, seed {S['seed']}, genome {glen} bp)"))


# ncu dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant classes, config 2, device-resident step
# (profiles/r2z_hot_kernels.txt, the `ncu --set full` capture of the final round-2 code): the `traffic` figure of the roofline object
TRAFFIC_CFG2 = {"group": 1938.9e6, "radix_pass": 2404.1e6, "cluster": 421.4e6, "keys": 1393.4e6, "pack": 284.3e6,
                "assemble": 366.0e6, "unpermute": 1079.7e6}      # pack / unpermute: mean over the class's launches


def headline(r, env, args, scaling, traffic):
    peak, peak_src = measured_peak_gbs()
    roof_all = roofline_table(r["prof"], r["steps"], r["ms"], peak)
    dom = max(roof_all, key=lambda k: roof_all[k]["ms_per_step"])
    d_ms, d_n, _ = r["prof"][dom]
    value = r["bases_per_step"] * r["steps"] / (r["ms"] * 1e-3) / 1e9
    e2e = r["bases_per_step"] * r["steps"] / (r["ms_e2e"] * 1e-3) / 1e9
    return {
        "metric": METRIC, "value": value, "unit": "Gbases/s", "n_gpus": env.world, "steps": r["steps"], "warmup": args.warmup,
        "ms_per_step": r["ms"] / r["steps"], "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": r["workload"], "rows_per_gpu": r["rows_per_gpu"],
                   "parallelism": f"block-per-gpu x{env.world}, no data-path collective" + (f", {r['blocks_per_gpu']} block(s) per GPU in sequence" if "blocks_per_gpu" in r else ""),
                   "l2": "inputs and every intermediate array exceed the 126 MB L2 many times over; no explicit flush"},
        "e2e": {"value": e2e, "unit": "Gbases/s", "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
                "ms_per_step": r["ms_e2e"] / r["steps"], "equals_device_result": r["e2e_equals_device"],
                **({"reads_only": {"value": r["bases_per_step"] * r["steps"] / (r["ms_e2e_reads"] * 1e-3) / 1e9, "unit": "Gbases/s",
                                   "ms_per_step": r["ms_e2e_reads"] / r["steps"], "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h_reads"],
                                   "equals_device_result": r["e2e_reads_equal"],
                                   "note": "the same host-buffer call with bwt/qs/lcp = NULL: only the smoothed reads come back (the eBWT files "
                                           "are the reference's intermediate files); informational, the headline e2e returns all five arrays"}}
                   if "ms_e2e_reads" in r else {})},
        "gpu_launches": r["launches"],
        "roofline": {"bound": roof_all[dom]["bound"], "kernel": f"{KNAME[dom]} ({dom})", "achieved": roof_all[dom]["achieved"], "peak": peak,
                     "unit": "GB/s", "frac": roof_all[dom]["frac"], "traffic": (traffic or {}).get(dom), "peak_source": peak_src,
                     "launches": d_n, "avg_launch_ms": d_ms / d_n if d_n else None, "share_of_step": d_ms / r["ms"],
                     "note": "achieved = DRAM-level algorithmic bytes of the class / its CUDA-event time (rank 0); an 'issue'-bound class is "
                             "limited by instruction issue and L2 gathers, its fraction of the HBM peak is reported for completeness"},
        "roofline_all": roof_all,
        "stages_ms_per_step": {k: v[0] / r["steps"] for k, v in r["prof"].items()},
        "clocks": r["clocks"], "stats": r["stats"],
    }


def secondary(r, with_e2e=True):
    out = {"workload": r["workload"], "value": r["bases_per_step"] * r["steps"] / (r["ms"] * 1e-3) / 1e9, "unit": "Gbases/s",
           "ms_per_step": r["ms"] / r["steps"], "steps": r["steps"], "gpu_launches": r["launches"], "stats": r["stats"]}
    if with_e2e and "ms_e2e" in r:
        out["e2e"] = {"value": r["bases_per_step"] * r["steps"] / (r["ms_e2e"] * 1e-3) / 1e9, "unit": "Gbases/s", "ms_per_step": r["ms_e2e"] / r["steps"],
                      "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"], "equals_device_result": r["e2e_equals_device"]}
    for k in ("checks_ok", "stage_ms_per_step", "kernel_class_ms_per_step", "rows", "wide_positions", "blocks_per_gpu"):
        if k in r:
            out[k] = r[k]
    if "prof" in r:
        out["stages_ms_per_step"] = {k: round(v[0] / r["steps"], 3) for k, v in r["prof"].items() if v[0] > 0}
    return out


def gpu_arm(args):
    env = Env()
    parts = set(args.parts.split(","))
    sec_steps = max(1, min(args.steps, 3))
    line = None
    if env.world == 1:
        r2 = bench_cfg2(env, args.steps, args.warmup)
        line = headline(r2, env, args, "weak", TRAFFIC_CFG2)
        if "cfg4" in parts:
            r4 = bench_cfg4(env, sec_steps, 1)
            line["config4_blocks"] = secondary(r4)
            line["config4_blocks"]["note"] = "all 8 blocks in sequence on one GPU: the N=1 point of the strong-scaling series that bench.py --gpus 2/4/8 reports as `value`"
    else:
        r4 = bench_cfg4(env, args.steps, args.warmup)
        line = headline(r4, env, args, "strong", None)
        line["checks_ok"] = r4["checks_ok"]
        if "cfg2" in parts:
            r2 = bench_cfg2(env, sec_steps, 1)
            line["weak_config2"] = secondary(r2)
            line["weak_config2"]["scaling"] = "weak"
    if "single" in parts:
        # two warm-up steps: the engine returns its slices in fresh torch tensors (8.5 GB per call at one rank) and the previous
        # result is still held when the next call allocates, so the allocator reaches its steady state after the second call
        rs = bench_single(env, sec_steps, 2)
        line["single_collection"] = secondary(rs, with_e2e=False)
        line["single_collection"]["collective"] = "one all_to_all_single (NCCL over NVLink) of the smoothed rows + an all_reduce of the statistics" if env.world > 1 else "none (one rank)"
    if env.rank == 0:
        cpu = None
        if env.world == 1 and not args.no_cpu_baseline:
            # bounded sample of the headline workload on this box's host cores (the reference arm times the full configuration)
            n = 100_000 // (int(os.environ.get("BFQ_BENCH_QUICK", "0")) or 1)
            W = CFG2
            if have_ref():
                t, parts_t = ref_cfg2(n, 1)
                cpu = {"value": n * W["read_len"] / t / 1e9, "unit": "Gbases/s", "cores": 1, "kind": "reference",
                       "sample": f"first {n} reads x {W['read_len']} bp of config 2 ({100.0 * n / W['nreads']:.0f} %, seed {W['seed']}), {t:.1f} s: gsufsort "
                                 f"{parts_t[0][0]:.1f} s + bfq_ext -T 1 {parts_t[0][1]:.1f} s (the thread count BFQzip_ext.py runs), unmodified reference binaries",
                       "host": host_info()}
            else:
                t = run_port_sample(40_000, W["read_len"], W["seed"], W["M"], W["B"])
                cpu = {"value": 40_000 * W["read_len"] / t / 1e9, "unit": "Gbases/s", "cores": 1, "kind": "port", "sample": "40000 reads of config 2; oracle C port"}
            rec = os.path.join(ROOT, "profiles", "r2_cpu_baseline_full.json")
            if os.path.exists(rec):
                with open(rec) as f:
                    cpu["recorded_full_config2"] = json.load(f)      # measured once on this pool's host cores, not in this run
        line["cpu_baseline"] = cpu
        _emit(line)
    if env.world > 1:
        env.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--parts", default="cfg2,cfg4,single", help="secondary measurements to include (the headline always runs)")
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
