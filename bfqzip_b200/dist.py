"""One collection over several GPUs: the single-eBWT mode of the pipeline (BFQzip_ext.py:173-196 runs eGap and
bfq_ext on the whole FASTQ; SURVEY.md section 8(e), second row).

Unlike the block mode (parallel.py, the BFQzip_parallel.py semantics) the result here is bit-identical to the
single-GPU pipeline on the same FASTQ -- one eBWT, one set of clusters -- and the ranks share the work:

  1. every rank holds the reads (they are replicated: 2 bytes per base, small next to the sort's working set);
  2. rank r sorts only the suffixes whose first four symbols fall in its key range and so builds rows
     [row_base, row_base + n_local) of the global eBWT / QS / LCP.  The ranges come from a histogram every rank
     computes identically: no splitter exchange, and equal keys never straddle ranks;
  3. the slice is smoothed locally.  That is exact for -k >= 4: a cluster needs LCP >= k between neighbouring rows
     and rows of different ranks share at most 3 symbols;
  4. the one exchange step: every smoothed row is a (text position, base, quality) triple; rank d rebuilds a
     contiguous range of reads, so the triples travel in ONE all-to-all (NCCL over NVLink) to the owner of their position;
  5. the owner sorts what it received back into text order and writes its reads.

Engines: ``GpuEngine`` drives the C ABI (bfq_dist_build / bfq_dist_place).  The orchestration takes any object with
the same methods, which is how the CPU tests run the exchange over gloo with a stand-in engine and no GPU.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

STAT_SUM_KEYS = ("clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified")


def read_bounds(off: np.ndarray, world: int) -> List[int]:
    """Reads [rb[d], rb[d+1]) are rebuilt by rank d: contiguous, balanced on text positions.
    rb[d] = first read g whose end in the text, off[g + 1] + g + 1, reaches total * d / world.  That end is monotone in g, so
    each bound is a bisection over the offsets: O(world * log nreads) -- the offsets of 50M reads are 400 MB, and building the
    array of ends on every call kept the GPU idle for a fifth of a second per step (BASELINE config 5 on 8 GPUs)."""
    nreads = len(off) - 1
    end = lambda g: int(off[g + 1]) + g + 1                    # first text position after read g
    total = end(nreads - 1) if nreads else 0
    rb = [0]
    for d in range(1, world):
        want = total * d // world
        lo, hi = 0, nreads                                     # first g in [0, nreads) with end(g) >= want, else nreads
        while lo < hi:
            mid = (lo + hi) // 2
            if end(mid) >= want:
                hi = mid
            else:
                lo = mid + 1
        rb.append(max(rb[-1], min(nreads, lo)))
    rb.append(nreads)
    return rb


def text_bounds(off: np.ndarray, rb: Sequence[int]) -> np.ndarray:
    """First text position (terminators included) of every rank's read range, plus the end."""
    return np.array([int(off[g]) + g for g in rb], dtype=np.uint64)


def slice_capacity(n_rows: int, world: int) -> int:
    """Rows a rank must be able to hold: the key ranges are cut on 4096 bins, so a slice can exceed n / world."""
    if world == 1:
        return n_rows + 64
    return n_rows // world + n_rows // (2 * world) + (1 << 16)


class GpuEngine:
    """The product path: C ABI calls on this rank's GPU.  reads are device tensors (u8 bases, u8 quals, i64 off)."""

    def __init__(self, M: int = 2, B: int = 0, device: int = 0, options: dict | None = None):
        import torch
        from . import lib
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.g = lib.BfqGpu(M, B, device)
        for k, v in (options or {}).items():
            self.g.set_option(k, int(v))

    def close(self):
        self.g.close()

    def upload(self, reads):
        t = self.torch
        bases, quals, off = reads
        d = (t.from_numpy(np.ascontiguousarray(bases)).to(self.dev), t.from_numpy(np.ascontiguousarray(quals)).to(self.dev),
             t.from_numpy(np.ascontiguousarray(off, np.uint64).view(np.int64)).to(self.dev))
        # the engine works on its own stream: the copies (torch's stream) must have landed before it reads them
        t.cuda.current_stream(self.dev).synchronize()
        return d

    def build(self, rank, world, dreads, nreads, total_bases, params, tb):
        t = self.torch
        d_b, d_q, d_off = dreads
        cap = slice_capacity(total_bases + nreads, world)
        u8 = lambda: t.empty(cap + 64, dtype=t.uint8, device=self.dev)
        bwt, qs, lcp = u8(), u8(), u8()
        keys = t.empty(cap + 16, dtype=t.int32, device=self.dev)
        vals = t.empty(cap + 16, dtype=t.int16, device=self.dev)
        n_local, row_base, counts, st = self.g.dist_build(d_b, d_q, d_off, nreads, total_bases, params, rank, world, cap, tb,
                                                          bwt, qs, lcp, keys, vals)
        return dict(n_local=n_local, row_base=row_base, bwt=bwt[:n_local], qs=qs[:n_local], lcp=lcp[:n_local],
                    keys=keys[:n_local], vals=vals[:n_local], counts=counts, stats=st)

    def place(self, keys, vals, count, text_lo, dreads, read_lo, read_hi, nbases):
        t = self.torch
        ob = t.empty(nbases + 64, dtype=t.uint8, device=self.dev); oq = t.empty_like(ob)
        self.g.dist_place(keys, vals, count, text_lo, dreads[2], read_lo, read_hi, ob, oq)
        return ob[:nbases], oq[:nbases]


def exchange_pairs(keys, vals, counts: np.ndarray, group=None):
    """The all-to-all of step 4.  keys i32 / vals i16 tensors sorted by destination, counts[d] = pairs for rank d."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    send = torch.tensor([int(x) for x in counts], dtype=torch.int64, device=keys.device)
    recv = torch.empty(world, dtype=torch.int64, device=keys.device)
    dist.all_to_all_single(recv, send, group=group)
    rc = [int(x) for x in recv.tolist()]
    sc = [int(x) for x in counts]
    rkeys = torch.empty(sum(rc) + 16, dtype=keys.dtype, device=keys.device)
    rvals = torch.empty(sum(rc) + 16, dtype=vals.dtype, device=vals.device)
    dist.all_to_all_single(rkeys[:sum(rc)], keys[:sum(sc)].contiguous(), rc, sc, group=group)
    # 16-bit integers are not a collective type (NCCL, gloo): the payload travels as bytes
    dist.all_to_all_single(rvals[:sum(rc)].view(torch.uint8), vals[:sum(sc)].contiguous().view(torch.uint8),
                           [2 * x for x in rc], [2 * x for x in sc], group=group)
    if rkeys.is_cuda:
        # the engine works on its own stream: the received pairs must have landed before it reads them
        torch.cuda.current_stream(rkeys.device).synchronize()
    return rkeys, rvals, sum(rc)


def run_single_collection(engine, reads, params, rank: int, world: int, group=None, dreads=None, timing: dict | None = None):
    """One rank's share.  Returns (slice dict, (out_bases, out_quals) of reads [read_lo, read_hi), (read_lo, read_hi), stats).
    timing (optional): filled with this rank's seconds per stage -- build (range selection, sort, group ordering, smoothing,
    pair emission: bfq_dist_build, which returns synchronised), exchange (the all-to-all), place (bfq_dist_place)."""
    import time
    import torch
    import torch.distributed as dist
    t0 = time.perf_counter()
    off = np.ascontiguousarray(reads[2], np.uint64)
    nreads = len(off) - 1
    total = int(off[-1])
    rb = read_bounds(off, world)
    tb = text_bounds(off, rb)
    if dreads is None:
        dreads = engine.upload(reads)
    sl = engine.build(rank, world, dreads, nreads, total, params, tb)
    t1 = time.perf_counter()
    if world > 1:
        rkeys, rvals, cnt = exchange_pairs(sl["keys"], sl["vals"], sl["counts"], group)
    else:
        rkeys, rvals, cnt = sl["keys"], sl["vals"], int(sl["counts"][0])
    t2 = time.perf_counter()
    lo, hi = rb[rank], rb[rank + 1]
    expect = int(tb[rank + 1] - tb[rank])
    if cnt != expect:
        raise RuntimeError(f"rank {rank}: received {cnt} rows for {expect} text positions")
    nb = int(off[hi] - off[lo])
    ob, oq = engine.place(rkeys, rvals, cnt, int(tb[rank]), dreads, lo, hi, nb)
    if timing is not None:
        t3 = time.perf_counter()
        for k, v in (("build", t1 - t0), ("exchange", t2 - t1), ("place", t3 - t2)):
            timing[k] = timing.get(k, 0.0) + v
    stats = dict(sl["stats"])
    if world > 1:
        v = torch.tensor([stats[k] for k in STAT_SUM_KEYS], dtype=torch.int64, device=rkeys.device)
        dist.all_reduce(v, group=group)
        for k, x in zip(STAT_SUM_KEYS, v.tolist()):
            stats[k] = int(x)
    return sl, (ob, oq), (lo, hi), stats


def emulate_ranks(engine, reads, params, world: int):
    """All ranks of the mode run one after the other on ONE engine, the exchange done in memory: what the single-GPU
    parity tests use.  Returns (bwt, qs, lcp, out_bases, out_quals, stats) of the whole collection as numpy arrays."""
    import torch
    off = np.ascontiguousarray(reads[2], np.uint64)
    nreads = len(off) - 1
    total = int(off[-1])
    rb = read_bounds(off, world)
    tb = text_bounds(off, rb)
    dreads = engine.upload(reads)
    slices = [engine.build(r, world, dreads, nreads, total, params, tb) for r in range(world)]
    order = np.argsort([s["row_base"] for s in slices], kind="stable")
    assert list(order) == list(range(world))
    cat = lambda k: np.concatenate([s[k].cpu().numpy() for s in slices])
    bwt, qs, lcp = cat("bwt"), cat("qs"), cat("lcp")
    stats = dict(slices[0]["stats"])
    for k in STAT_SUM_KEYS:
        stats[k] = sum(s["stats"][k] for s in slices)
    obs, oqs = [], []
    for d in range(world):
        ks, vs = [], []
        for s in slices:
            c = [int(x) for x in s["counts"]]
            a = sum(c[:d])
            ks.append(s["keys"][a:a + c[d]]); vs.append(s["vals"][a:a + c[d]])
        pad_k = torch.zeros(16, dtype=ks[0].dtype, device=ks[0].device); pad_v = torch.zeros(16, dtype=vs[0].dtype, device=vs[0].device)
        rk = torch.cat(ks + [pad_k]); rv = torch.cat(vs + [pad_v])
        cnt = rk.numel() - 16
        assert cnt == int(tb[d + 1] - tb[d]), (d, cnt, int(tb[d + 1] - tb[d]))
        lo, hi = rb[d], rb[d + 1]
        # torch.cat runs on torch's stream, the engine on its own: at 10^8 pairs the engine would otherwise start on a
        # half-written list (exchange_pairs synchronises the same way after the all-to-all)
        torch.cuda.current_stream(rk.device).synchronize()
        ob, oq = engine.place(rk, rv, cnt, int(tb[d]), dreads, lo, hi, int(off[hi] - off[lo]))
        obs.append(ob.cpu().numpy()); oqs.append(oq.cpu().numpy())
    return bwt, qs, lcp, np.concatenate(obs), np.concatenate(oqs), stats
