"""Block-partitioned mode (bfqzip_b200/parallel.py): partition / merge semantics of
BFQzip_parallel.py, and the N>1 scheduling over torch.distributed (gloo, world_size 2, on CPU)."""
import os
import socket
import sys

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import parallel, synth


def ref_split(nreads, threads):
    """BFQzip_parallel.py:288-323 restated on read indices (the loop that fills the block files)."""
    size_block = nreads if threads == 0 else nreads // threads
    num_blocks = nreads // size_block
    blocks, nxt = [], 0
    for i in range(num_blocks):
        cnt = 0
        start = nxt
        while nxt < nreads:
            nxt += 1; cnt += 1
            if cnt == size_block and i + 1 != num_blocks:
                break
        blocks.append((start, nxt))
    return blocks


@pytest.mark.parametrize("nreads,threads", [(10, 4), (100, 8), (101, 8), (7, 7), (8, 3), (5, 0), (1, 1), (1000, 16)])
def test_split_matches_reference_loop(nreads, threads):
    got = parallel.split_blocks(nreads, threads)
    assert got == ref_split(nreads, threads)
    assert got[0][0] == 0 and got[-1][1] == nreads
    assert all(a[1] == b[0] for a, b in zip(got, got[1:]))


def test_split_ten_reads_four_threads_gives_five_blocks():
    assert len(parallel.split_blocks(10, 4)) == 5


def test_blocks_of_rank_is_a_contiguous_partition():
    for nb in (1, 5, 8, 13):
        for world in (1, 2, 4, 8):
            allb = [b for r in range(world) for b in parallel.blocks_of_rank(nb, r, world)]
            assert allb == list(range(nb))
    assert [len(parallel.blocks_of_rank(8, r, 4)) for r in range(4)] == [2, 2, 2, 2]


def _tag(reads):
    """A stand-in per-block transform that depends on the WHOLE block (like an eBWT does)."""
    b, q, off = reads
    k = np.uint8(len(off) - 1)
    return b.copy(), ((q.astype(np.uint16) + k) % 94 + 33).astype(np.uint8), off.copy()


def test_paired_blocks_append_mate2_and_deinterleave():
    r1 = synth.make_reads(41, 30, 1, var_len=True)
    r2 = synth.make_reads(41, 25, 2, var_len=True)
    blocks = parallel.make_blocks(r1, 4, r2)
    cuts = parallel.split_blocks(41, 4)
    assert len(blocks) == len(cuts)
    for (blk, n1), (s, e) in zip(blocks, cuts):
        assert n1 == e - s and len(blk[2]) - 1 == 2 * (e - s)
        a = parallel.slice_reads(r1, s, e); b = parallel.slice_reads(r2, s, e)
        assert np.array_equal(blk[0], np.concatenate([a[0], b[0]]))
    m1, m2 = parallel.run_blocks(r1, 4, lambda x: (x[0], x[1], x[2]), r2)
    for got, exp in ((m1, r1), (m2, r2)):
        assert all(np.array_equal(g, e) for g, e in zip(got, exp))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        r1 = synth.make_reads(203, 40, 5, var_len=True)
        r2 = synth.make_reads(203, 40, 6, var_len=True)
        single = parallel.run_blocks(r1, 8, _tag, rank=rank, world=world)
        paired = parallel.run_blocks(r1, 8, _tag, r2, rank=rank, world=world)
        # the per-block-file path (what real sizes use) must give the same merged collections as the pickled gather
        spill = os.path.join(os.environ["BFQ_TEST_SPILL"], "blocks")
        single_f = parallel.run_blocks(r1, 8, _tag, rank=rank, world=world, spill_dir=spill + "_s")
        paired_f = parallel.run_blocks(r1, 8, _tag, r2, rank=rank, world=world, spill_dir=spill + "_p")
        if rank == 0:
            for a, b in zip(single, single_f):
                assert np.array_equal(a, np.asarray(b))
            for ma, mb in zip(paired, paired_f):
                for a, b in zip(ma, mb):
                    assert np.array_equal(a, np.asarray(b))
            assert len(os.listdir(spill + "_p")) == 3 * 8
            q.put((single, paired))
        else:
            assert single is None and paired is None and single_f is None and paired_f is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_equal_one_rank(tmp_path, monkeypatch):
    """The result depends on the block count, not on the rank count -- through both gather paths."""
    import torch.multiprocessing as mp
    monkeypatch.setenv("BFQ_TEST_SPILL", str(tmp_path))
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    single2, paired2 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    r1 = synth.make_reads(203, 40, 5, var_len=True)
    r2 = synth.make_reads(203, 40, 6, var_len=True)
    single1 = parallel.run_blocks(r1, 8, _tag)
    paired1 = parallel.run_blocks(r1, 8, _tag, r2)
    assert all(np.array_equal(a, b) for a, b in zip(single1, single2))
    for m in (0, 1):
        assert all(np.array_equal(a, b) for a, b in zip(paired1[m], paired2[m]))


@pytest.mark.gpu
def test_gpu_blocks_equal_oracle_blocks():
    """8 blocks through the CUDA pipeline == 8 independent oracle runs, merged in block order."""
    from bfqzip_b200 import lib
    r1 = synth.make_reads(4000, 60, 11)
    r2 = synth.make_reads(4000, 60, 12)
    g = lib.BfqGpu(3, 0, 0)
    P = lib.default_params(eos=0)

    def gpu_block(reads):
        _, (ob, oq), _ = g.pipeline(reads[0], reads[1], reads[2], P, want_files=False)
        return ob, oq, reads[2]

    def oracle_block(reads):
        b, q, l = H.oracle_build_ebwt(reads[0], reads[1], reads[2], 0)
        ob, oq, off, st, _, _, err = H.oracle_smooth_rebuild(b, q, l, H.make_params(M=3, B=0, eos=0))
        assert err == 0
        return ob, oq, off

    got = parallel.run_blocks(r1, 8, gpu_block, r2)
    exp = parallel.run_blocks(r1, 8, oracle_block, r2)
    for m in (0, 1):
        assert all(np.array_equal(a, b) for a, b in zip(got[m], exp[m]))
    g.close()


def test_counter_based_generator_blocks_equal_the_split_of_the_whole_collection():
    """bench.py generates every config-4 block on the GPU that processes it (synth.hash_paired_block_torch): the block must
    be exactly what parallel.make_blocks cuts out of the whole paired collection (split_fastq / split_fastq_2 semantics),
    and the torch version must give the bytes of the numpy version."""
    import torch
    npairs, L, seed, frag, nb = 4100, 40, 1003, 90, 8
    r1, r2 = synth.hash_paired(npairs, L, seed, coverage=40.0, fragment=frag)
    g = synth.hash_genome_len(2 * npairs, L, 40.0, frag)
    blocks = parallel.make_blocks(r1, nb, r2)
    cuts = parallel.split_blocks(npairs, nb)
    assert len(blocks) == len(cuts) == nb
    for (reads, n1), (lo, hi) in zip(blocks, cuts):
        b, q, off = synth.hash_paired_block_torch(torch, torch.device("cpu"), npairs, L, seed, g, frag, lo, hi)
        assert n1 == hi - lo
        assert np.array_equal(reads[0], b.numpy()) and np.array_equal(reads[1], q.numpy()) and np.array_equal(reads[2], off)
    # the model: every read comes from the genome (mate 2 reverse-complemented), ~1 % errors, N at offset 7 of every 1000th read
    assert set(np.unique(r1[0]).tolist()) == set(b"ACGNT") and r1[0][7] == ord("N") and r1[1][7] == ord("#")
    low = (r1[1] < 33 + 20).mean()
    assert 0.005 < low < 0.02
    # chunking on the device does not change the result
    a = synth.hash_reads_torch(torch, torch.device("cpu"), 1000, 30, 5, 2000, chunk_reads=64)
    h = synth.hash_reads(1000, 30, 5, 2000)
    assert np.array_equal(a[0].numpy(), h[0]) and np.array_equal(a[1].numpy(), h[1])
