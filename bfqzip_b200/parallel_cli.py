"""Command-line launcher with the flag surface of BFQzip_parallel.py (reference file:line 27-37), driving read blocks
onto GPUs instead of CPU threads.

    python -m bfqzip_b200.parallel_cli reads.fastq -t 8 -0 [-o OUT] [-T K] [-Q CHAR] [-H] [--reorder 1 --seed S]
    python -m bfqzip_b200.parallel_cli reads_1.fastq reads_2.fastq -p -t 8 -0
    torchrun --nproc-per-node N ... -m bfqzip_b200.parallel_cli ...        # one rank per GPU, blocks dealt to ranks

Same partition and merge as the reference: ``-t n`` cuts ``size_block = R // n`` reads per block (the last block takes the
remainder), paired mode appends mate-2 block i to mate-1 block i, outputs are concatenated in block order and paired output
is de-interleaved (BFQzip_parallel.py:288-360, 153-177).  Output names follow :137-150: ``IN.cat.EXT`` (or ``OUT.EXT``;
``OUT_1.EXT`` / ``OUT_2.EXT`` in paired mode).  The result depends on the number of blocks, never on the number of GPUs.

Differences, all stated: ``--reorder 1`` takes ``--seed`` (the reference's randomFASTQ.py is unseeded); ``--reorder 2``
(SPRING) and the entropy coders of step 5 stay on the host and are not run here -- without ``-0`` the launcher stops after
the merge and says so; ``-M``/``-B`` pick the compile-time mode of the library (the reference rebuilds its binaries).
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np

from . import fastq, parallel


def build_parser():
    ap = argparse.ArgumentParser(prog="bfqzip_b200.parallel_cli", description=__doc__.split("\n\n")[0])
    ap.add_argument("input", nargs="+", help="input file name(s)")
    ap.add_argument("-o", "--out", default="", help="output base name (def. input base name)")
    ap.add_argument("-T", "--mcl", default="", help="minimum context length")
    ap.add_argument("-Q", "--rv", default="", help="constant replacement value")
    ap.add_argument("-H", "--headers", action="store_true", help="store original headers")
    ap.add_argument("-0", "--m0", action="store_true", help="mode 0: do not compress")
    ap.add_argument("-p", "--paired", action="store_true", help="paired end mode")
    ap.add_argument("-t", "--threads", default=0, type=int, help="number of blocks (0: one block)")
    ap.add_argument("--reorder", default=0, type=int, help="reorder reads (0: no reorder, 1: random)")
    ap.add_argument("--seed", default=0, type=int, help="seed of --reorder 1")
    ap.add_argument("-c", "--check", action="store_true", help="check that the FASTQ is valid")
    ap.add_argument("-v", default=0, type=int, help="verbose")
    ap.add_argument("-M", default=2, type=int, help="quality smoothing mode of the library (make M=)")
    ap.add_argument("-B", default=0, type=int, help="Illumina 8-level binning (make B=)")
    return ap


def output_names(args):
    """BFQzip_parallel.py:137-150."""
    base, ext = os.path.splitext(args.input[0])
    if not args.out:
        if args.paired:
            base2, ext2 = os.path.splitext(args.input[1])
            return [base + ".cat" + ext, base2 + ".cat" + ext2]
        return [base + ".cat" + ext]
    if args.paired:
        return [args.out + "_1" + ext, args.out + "_2" + ext]
    return [args.out + ext]


def reorder_reads(reads, headers, seed):
    """A seeded random permutation of the records (what ``--reorder 1`` does with randomFASTQ.py)."""
    b, q, off = reads
    n = len(off) - 1
    perm = np.random.default_rng(seed).permutation(n)
    o = off.astype(np.int64)
    lens = (o[1:] - o[:-1])[perm]
    noff = np.zeros(n + 1, np.uint64)
    np.cumsum(lens, out=noff[1:].view(np.int64))
    idx = np.concatenate([np.arange(o[g], o[g + 1]) for g in perm]) if n else np.empty(0, np.int64)
    return (b[idx], q[idx], noff), [headers[g] for g in perm]


def load(path, check):
    b, q, off, headers = fastq.read_fastq(path)
    if check and not all(h.startswith(b"@") for h in headers):
        raise SystemExit(f"Invalid FASTQ file: {path}")
    return (b, q, off), headers


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.paired and len(args.input) != 2:
        print("=== ERROR ===\npaired end mode")
        return 1
    if args.reorder == 2:
        print("--reorder 2 (SPRING) runs on the host and is not part of this launcher")
        return 1
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")                  # host-side gather of finished blocks only: no data-path collective
        group = dist.group.WORLD
    from . import lib
    g = lib.BfqGpu(args.M, args.B, local)
    P = lib.default_params(eos=0)
    if args.mcl:
        P.K = int(args.mcl)
    if args.rv:
        P.default_q = ord(args.rv)
    t0 = time.time()
    reads1, heads1 = load(args.input[0], args.check)
    reads2, heads2 = (load(args.input[1], args.check) if args.paired else (None, None))
    if args.reorder == 1:
        reads1, heads1 = reorder_reads(reads1, heads1, args.seed)
        if args.paired:
            reads2, heads2 = reorder_reads(reads2, heads2, args.seed)      # the same permutation keeps the mates aligned
    if rank == 0:
        print("Threads =", args.threads)

    def gpu_block(block):
        _, (ob, oq), _ = g.pipeline(block[0], block[1], block[2], P, want_files=False)
        return ob, oq, block[2]

    # several ranks: finished blocks go through per-block files next to the output (removed afterwards), not through pickles
    spill = None
    if world > 1:
        spill = os.path.join(os.path.dirname(os.path.abspath(output_names(args)[0])), ".bfq_blocks_" + os.environ.get("MASTER_PORT", "0"))
    merged = parallel.run_blocks(reads1, args.threads, gpu_block, reads2, rank=rank, world=world, gather_group=group, spill_dir=spill)
    g.close()
    if rank == 0:
        names = output_names(args)
        outs = merged if args.paired else (merged,)
        for name, reads, heads in zip(names, outs, (heads1, heads2)):
            fastq.write_fastq(name, reads[0], reads[1], reads[2], heads if args.headers else None)
            print(name)
        print("Total time: {0:.4f}".format(time.time() - t0))
        if not args.m0:
            print("step 5 (PPMd / BSC entropy coding) stays on the host and is not run by this launcher; "
                  "pass the merged files to the reference's compressors")
    if world > 1:
        import shutil
        import torch.distributed as dist
        dist.barrier()
        if rank == 0 and spill:
            shutil.rmtree(spill, ignore_errors=True)
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
