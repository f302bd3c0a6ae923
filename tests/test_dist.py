"""Host logic of the single-collection mode (bfqzip_b200/dist.py) on CPU: ownership bounds, and the exchange step
over torch.distributed (gloo, world_size 2) with a stand-in engine built from the oracle in place of the GPU."""
import os
import socket

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import dist, synth


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_read_bounds_partition_the_text(world):
    _, _, off = synth.make_reads(57, 40, 3, var_len=True)
    rb = dist.read_bounds(off, world)
    assert rb[0] == 0 and rb[-1] == 57 and len(rb) == world + 1
    assert all(a <= b for a, b in zip(rb, rb[1:]))
    tb = dist.text_bounds(off, rb)
    assert int(tb[0]) == 0 and int(tb[-1]) == int(off[-1]) + 57
    sizes = np.diff(tb.astype(np.int64))
    assert sizes.max() - sizes.min() <= 2 * 41 or world == 1      # balanced up to one read


def test_read_bounds_more_ranks_than_reads():
    off = np.array([0, 10, 20], np.uint64)
    rb = dist.read_bounds(off, 8)
    assert rb[0] == 0 and rb[-1] == 2 and all(a <= b for a, b in zip(rb, rb[1:]))


def test_slice_capacity_covers_an_even_split():
    for n in (10, 10**6, 10**9):
        for w in (1, 2, 8):
            assert dist.slice_capacity(n, w) >= (n + w - 1) // w


class OracleEngine:
    """Stand-in for GpuEngine: whole-collection oracle results cut into row slices at rows with LCP < 4."""

    def __init__(self, M, B, eos):
        self.M, self.B, self.eos = M, B, eos

    def upload(self, reads):
        return reads

    def _whole(self, reads):
        import torch
        bases, quals, off = reads
        nreads = len(off) - 1
        # text with per-read terminators ordered by read index, suffix array by plain sorting (small inputs only)
        text = []
        for g in range(nreads):
            text += [int(x) + nreads for x in bases[int(off[g]):int(off[g + 1])]] + [g]
        n = len(text)
        sa = sorted(range(n), key=lambda p: text[p:])
        bwt, qs, lcp = H.oracle_build_ebwt(bases, quals, off, self.eos)
        sq, ob, st, err = H.oracle_smooth(bwt, qs, lcp, H.make_params(M=self.M, B=self.B, eos=self.eos))
        assert err == 0
        keys = np.array([(p - 1) if p else n - 1 for p in sa], np.int64)
        vals = ob.astype(np.int64) | (sq.astype(np.int64) << 8)
        return bwt, qs, lcp, keys, vals, st, torch

    def build(self, rank, world, dreads, nreads, total, params, tb):
        bwt, qs, lcp, keys, vals, st, torch = self._whole(dreads)
        n = len(bwt)
        cuts = [0]
        for r in range(1, world):
            i = max(cuts[-1], n * r // world)
            while i < n and lcp[i] >= 4:
                i += 1
            cuts.append(i)
        cuts.append(n)
        a, b = cuts[rank], cuts[rank + 1]
        k, v = keys[a:b], vals[a:b]
        o = np.argsort(k, kind="stable")
        k, v = k[o], v[o]
        counts = np.array([np.count_nonzero((k >= int(tb[d])) & (k < int(tb[d + 1]))) for d in range(world)], np.uint64)
        dest = np.searchsorted(tb.astype(np.int64), k, side="right") - 1
        k = k - tb.astype(np.int64)[dest]                    # positions travel relative to the destination's first one
        stats = {kk: (vv if rank == 0 else 0) for kk, vv in st.items()}        # the counters are summed over ranks
        return dict(n_local=b - a, row_base=a, bwt=bwt[a:b], qs=qs[a:b], lcp=lcp[a:b],
                    keys=torch.from_numpy(k.astype(np.int32)), vals=torch.from_numpy(v.astype(np.int16)), counts=counts, stats=stats)

    def place(self, keys, vals, count, text_lo, dreads, read_lo, read_hi, nbases):
        import torch
        off = dreads[2]
        k = keys[:count].numpy().astype(np.int64)
        v = vals[:count].numpy().astype(np.int64) & 0xFFFF
        assert np.array_equal(np.sort(k), np.arange(count))
        t = np.empty(count, np.int64); t[k] = v
        ob, oq = [], []
        for g in range(read_lo, read_hi):
            s = int(off[g]) + g - text_lo
            seg = t[s:s + int(off[g + 1] - off[g])]
            ob.append((seg & 0xFF).astype(np.uint8)); oq.append((seg >> 8).astype(np.uint8))
        cat = lambda x: np.concatenate(x) if x else np.empty(0, np.uint8)
        return torch.from_numpy(cat(ob)), torch.from_numpy(cat(oq))


def _reads():
    return synth.make_reads(60, 30, 21, var_len=True)


def _worker(rank, world, port, q):
    import torch.distributed as td
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sl, (ob, oq), (lo, hi), st = dist.run_single_collection(OracleEngine(2, 0, 0), _reads(), None, rank, world)
        q.put((rank, sl["row_base"], np.asarray(sl["bwt"]), ob.numpy(), oq.numpy(), lo, hi, st))
        td.barrier()
    finally:
        td.destroy_process_group()


def test_two_ranks_gloo_equal_whole_collection():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=180) for _ in range(2)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    bases, quals, off = _reads()
    bwt, qs, lcp = H.oracle_build_ebwt(bases, quals, off, 0)
    eb, eq, eoff, est, _, _, err = H.oracle_smooth_rebuild(bwt, qs, lcp, H.make_params(M=2, B=0, eos=0))
    assert got[0][1] == 0 and got[1][1] == len(got[0][2])
    assert np.array_equal(np.concatenate([got[0][2], got[1][2]]), bwt)
    assert got[0][5] == 0 and got[0][6] == got[1][5] and got[1][6] == len(off) - 1
    assert np.array_equal(np.concatenate([got[0][3], got[1][3]]), eb)
    assert np.array_equal(np.concatenate([got[0][4], got[1][4]]), eq)
    for r in (0, 1):                                     # all-reduced counters are the whole collection's on every rank
        assert all(got[r][7][k] == est[k] for k in dist.STAT_SUM_KEYS)
