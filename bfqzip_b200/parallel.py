"""Block-partitioned multi-GPU mode: the partition / merge semantics of BFQzip_parallel.py with one
read block per GPU at a time and no data-path collective.

Mirrors (reference file:line):
  * ``split_fastq``   BFQzip_parallel.py:288-323  size_block = R // t, num_blocks = R // size_block,
                      the last block takes the remainder (so 10 reads, t=4 -> 5 blocks of 2);
  * ``split_fastq_2`` BFQzip_parallel.py:325-360  paired mode APPENDS mate-file block i to block
                      file i: block i = mates-1[blk i] ++ mates-2[blk i], one collection;
  * merge             BFQzip_parallel.py:153-177  blocks concatenated in block order; paired
                      outputs de-interleaved by the recorded mate-1 count of every block.
The result depends on the number of BLOCKS, never on the number of GPUs: blocks are dealt to
ranks in contiguous runs and every block is an independent collection with its own eBWT.

Ranks are ``torch.distributed`` processes (one per GPU).  Finished blocks reach rank 0 through per-block
files like the reference's ``cat`` (``spill_dir``), or -- small inputs only -- through a gloo ``gather_object``.  Nothing here touches the GPU itself: ``process_block`` does.
"""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import numpy as np

Reads = Tuple[np.ndarray, np.ndarray, np.ndarray]      # bases u8, quals u8, off u64[nreads+1]


def split_blocks(nreads: int, threads: int) -> List[Tuple[int, int]]:
    """Read ranges [start, end) of every block, exactly as split_fastq cuts them."""
    if nreads <= 0:
        return []
    size_block = nreads if threads <= 0 else nreads // threads
    if size_block == 0:
        # the reference divides by zero here (num_reads // 0); refuse instead of guessing
        raise ValueError(f"{threads} blocks requested for {nreads} reads")
    num_blocks = nreads // size_block
    out = []
    for i in range(num_blocks):
        s = i * size_block
        e = nreads if i + 1 == num_blocks else s + size_block
        out.append((s, e))
    return out


def slice_reads(reads: Reads, s: int, e: int) -> Reads:
    bases, quals, off = reads
    a, b = int(off[s]), int(off[e])
    return bases[a:b], quals[a:b], (off[s:e + 1] - off[s]).astype(np.uint64)


def concat_reads(parts: Sequence[Reads]) -> Reads:
    bases = np.concatenate([p[0] for p in parts]) if parts else np.empty(0, np.uint8)
    quals = np.concatenate([p[1] for p in parts]) if parts else np.empty(0, np.uint8)
    offs = [np.zeros(1, np.uint64)]
    base = np.uint64(0)
    for p in parts:
        offs.append(p[2][1:].astype(np.uint64) + base)
        base = base + np.uint64(p[2][-1])
    return bases, quals, np.concatenate(offs)


def make_blocks(reads1: Reads, threads: int, reads2: Reads | None = None):
    """Returns [(block reads, number of mate-1 reads in the block)] in block order."""
    n1 = len(reads1[2]) - 1
    cuts1 = split_blocks(n1, threads)
    if reads2 is None:
        return [(slice_reads(reads1, s, e), e - s) for s, e in cuts1]
    n2 = len(reads2[2]) - 1
    cuts2 = split_blocks(n2, threads)
    if len(cuts2) != len(cuts1):
        # split_fastq_2 would append to block files that split_fastq never made
        raise ValueError("mate files split into different numbers of blocks")
    blocks = []
    for (s1, e1), (s2, e2) in zip(cuts1, cuts2):
        blocks.append((concat_reads([slice_reads(reads1, s1, e1), slice_reads(reads2, s2, e2)]), e1 - s1))
    return blocks


def blocks_of_rank(num_blocks: int, rank: int, world: int) -> range:
    """Contiguous run of blocks for a rank (8 blocks on 1/2/4/8 GPUs -> 8/4/2/1 blocks each)."""
    lo = (num_blocks * rank) // world
    hi = (num_blocks * (rank + 1)) // world
    return range(lo, hi)


def merge_blocks(done: Sequence[Tuple[Reads, int]], paired: bool):
    """Block outputs, in block order -> one collection (or the two mate collections)."""
    if not paired:
        return concat_reads([r for r, _ in done])
    m1 = [slice_reads(r, 0, k) for r, k in done]
    m2 = [slice_reads(r, k, len(r[2]) - 1) for r, k in done]
    return concat_reads(m1), concat_reads(m2)


def _spill_paths(spill_dir: str, b: int):
    import os
    return [os.path.join(spill_dir, f"block_{b:05d}.{e}") for e in ("bases", "quals", "off")]


def run_blocks(reads1: Reads, threads: int, process_block: Callable[[Reads], Reads],
               reads2: Reads | None = None, rank: int = 0, world: int = 1, gather_group=None, spill_dir: str | None = None):
    """Every rank processes its run of blocks; rank 0 returns the merged result, others None.

    ``process_block`` maps one block (bases, quals, off) to (out_bases, out_quals, out_off); on a GPU
    rank it is ``BfqGpu.pipeline`` of that rank's context.  Finished blocks reach rank 0 in one of two ways:

    * ``spill_dir`` (a directory every rank of the node sees): each rank writes its blocks as three raw files per block and
      rank 0 memory-maps them after a barrier on ``gather_group`` -- what the reference does with its per-block output files
      and ``cat`` (BFQzip_parallel.py:153-177).  This is the path for real sizes: a config-4 run moves 12 GB of reads.
    * no ``spill_dir``: ``gather_object`` on ``gather_group`` (pickled arrays through gloo; fine for small inputs and tests).
    """
    blocks = make_blocks(reads1, threads, reads2)
    mine = blocks_of_rank(len(blocks), rank, world)
    if spill_dir is not None and world > 1:
        import os
        import torch.distributed as dist
        os.makedirs(spill_dir, exist_ok=True)
        for b in mine:
            out = process_block(blocks[b][0])
            for path, arr, dt in zip(_spill_paths(spill_dir, b), out, (np.uint8, np.uint8, np.uint64)):
                np.ascontiguousarray(arr, dt).tofile(path)
        dist.barrier(group=gather_group)                  # every block file is complete
        if rank != 0:
            return None
        everything = []
        for b in range(len(blocks)):
            pb, pq, po = _spill_paths(spill_dir, b)
            off = np.fromfile(po, np.uint64)
            rd = lambda path: np.memmap(path, np.uint8, "r") if os.path.getsize(path) else np.empty(0, np.uint8)
            everything.append((b, (rd(pb), rd(pq), off)))
        return merge_blocks([(out, blocks[b][1]) for b, out in everything], reads2 is not None)
    local = [(b, process_block(blocks[b][0])) for b in mine]
    if world == 1:
        everything = local
    else:
        import torch.distributed as dist
        gathered = [None] * world if rank == 0 else None
        dist.gather_object(local, gathered, dst=0, group=gather_group)
        if rank != 0:
            return None
        everything = [x for part in gathered for x in part]
    everything.sort(key=lambda t: t[0])
    assert [b for b, _ in everything] == list(range(len(blocks)))
    return merge_blocks([(out, blocks[b][1]) for b, out in everything], reads2 is not None)
