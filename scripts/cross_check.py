#!/usr/bin/env python
"""A small pass over every product path whose results are checked against each other: fused pipeline with both part-C
paths, files-in step 3, deferred-group levels, the single-collection slices (32- and 64-bit positions) and the FASTQ text
entry points.  (Written to run under compute-sanitizer; the tool is closed on the GPU pool, so it runs plain.)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
from bfqzip_b200 import dist, fastq, lib, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
reads = synth.make_reads(n, 80, 3, var_len=True)
b, q, off = reads
P = lib.default_params(eos=0)
g = lib.BfqGpu(2, 0, 0)
outs = []
for opts in ({"rebuild": 1}, {"rebuild": 2}, {"rebuild": 2, "group_cap": 8}):
    for k, v in opts.items():
        g.set_option(k, v)
    (bwt, qs, lcp), (ob, oq), st = g.pipeline(b, q, off, P)
    outs.append((bwt, qs, lcp, ob, oq, st))
    g.set_option("group_cap", 1 << 20)
for o in outs[1:]:
    assert all(np.array_equal(x, y) for x, y in zip(o[:5], outs[0][:5])) and o[5] == outs[0][5]
bwt, qs, lcp, ob, oq, st = outs[0]
sb, sq, soff, st2 = g.smooth_rebuild(bwt, qs, lcp, P)
assert np.array_equal(sb, ob) and np.array_equal(sq, oq) and st2 == st
files, text, st3 = g.pipeline_fastq(fastq.format_fastq(b, q, off), P)
assert text == fastq.format_fastq(ob, oq, off) and np.array_equal(files[0], bwt)
g.close()
for opt in ({}, {"wide_positions": 1}):
    eng = dist.GpuEngine(2, 0, 0, opt)
    dbwt, dqs, dlcp, dob, doq, dst = dist.emulate_ranks(eng, reads, P, 3)
    eng.close()
    assert np.array_equal(dbwt, bwt) and np.array_equal(dlcp, lcp) and np.array_equal(dob, ob) and np.array_equal(doq, oq)
print("cross_check: all paths agree")
