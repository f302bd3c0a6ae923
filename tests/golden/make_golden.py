"""Generates tests/golden/*.npz by running the UNMODIFIED reference binaries (oracle/_ref,
built by oracle/Makefile from /root/reference) on small inputs.  Run in the build container:

    make -C oracle ref && python tests/golden/make_golden.py

Each .npz holds the input reads, the reference's .bwt/.bwt.qs/.1.lcp bytes (gsufsort, EOS 35),
and for every (M,B) mode the reference's rebuilt reads + its printed statistics
(bfq_ext -l 250 -s 35 -m 5, K=16 default).  The GPU box has no /root/reference: tests only
read these files.
"""
import os, sys, tempfile
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers as H
from bfqzip_b200 import fastq, synth

MODES = [(M, B) for M in range(4) for B in range(2)]
STAT_KEYS = ["clusters", "discarded", "ambiguous", "processed", "bases_inside", "qs_smoothed", "modified"]


def make(name, bases, quals, off, headers=None):
    wd = tempfile.mkdtemp()
    fq = os.path.join(wd, "in.fastq")
    fastq.write_fastq(fq, bases, quals, off, headers)
    bwt, qs, lcp = H.ref_gsufsort(fq, wd)
    d = dict(bases=bases, quals=quals, off=off, bwt=bwt, qs=qs, lcp=lcp)
    if headers is not None:
        d["headers"] = np.frombuffer(b"\n".join(headers), dtype=np.uint8)
    for M, B in MODES:
        out, st = H.ref_bfq_ext(bwt, qs, lcp, M=M, B=B, headers=headers)
        ob, oq, ooff, _ = fastq.parse_fastq_bytes(out)
        assert (ooff == off).all()
        d[f"out_bases_M{M}_B{B}"] = ob
        d[f"out_quals_M{M}_B{B}"] = oq
        d[f"stats_M{M}_B{B}"] = np.array([st[k] for k in STAT_KEYS], dtype=np.uint64)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **d)
    print(name, "n =", len(bwt), "reads =", len(off) - 1)


if __name__ == "__main__":
    b, q, o, h = fastq.read_fastq("/root/reference/example/reads.fastq")
    make("example", b, q, o, h)
    b, q, o = synth.make_reads(1500, 64, 101, coverage=50)
    make("syn_fixed", b, q, o)
    # variable lengths + exact duplicate reads (tie-break by read index, SURVEY 5.1 rule 3)
    b, q, o = synth.make_reads(1200, 72, 202, coverage=60, var_len=True)
    lens = np.diff(o.astype(np.int64))
    dup = [5, 17, 5, 300]
    bb = [b] + [b[int(o[i]):int(o[i + 1])] for i in dup]
    qq = [q] + [q[int(o[i]):int(o[i + 1])] for i in dup]
    lens2 = np.concatenate([lens, lens[dup]])
    o2 = np.zeros(len(lens2) + 1, np.uint64); o2[1:] = np.cumsum(lens2)
    make("syn_var_dup", np.concatenate(bb), np.concatenate(qq), o2)
    # short reads, shorter than K in part: many tiny clusters / none
    b, q, o = synth.make_reads(400, 24, 303, coverage=30, var_len=True)
    make("syn_short", b, q, o)
