"""The BFQzip_parallel.py-compatible launcher (bfqzip_b200/parallel_cli.py): names, reordering, and (GPU) a whole run
against independent oracle runs of every block."""
import argparse
import os

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import fastq, parallel, parallel_cli, synth


def ns(**kw):
    d = dict(input=["a/reads.fastq"], out="", paired=False)
    d.update(kw)
    return argparse.Namespace(**d)


def test_output_names_follow_the_reference():
    assert parallel_cli.output_names(ns()) == ["a/reads.cat.fastq"]
    assert parallel_cli.output_names(ns(out="x/y")) == ["x/y.fastq"]
    assert parallel_cli.output_names(ns(input=["r_1.fq", "r_2.fq"], paired=True)) == ["r_1.cat.fq", "r_2.cat.fq"]
    assert parallel_cli.output_names(ns(input=["r_1.fq", "r_2.fq"], paired=True, out="o")) == ["o_1.fq", "o_2.fq"]


def test_reorder_is_a_seeded_permutation_of_records():
    reads = synth.make_reads(50, 30, 4, var_len=True)
    heads = [b"@%d" % i for i in range(50)]
    (b, q, off), h = parallel_cli.reorder_reads(reads, heads, 7)
    (b2, q2, off2), h2 = parallel_cli.reorder_reads(reads, heads, 7)
    assert np.array_equal(b, b2) and h == h2 and h != heads
    o = reads[2].astype(np.int64)
    recs = sorted((bytes(reads[0][o[i]:o[i + 1]]), bytes(reads[1][o[i]:o[i + 1]])) for i in range(50))
    p = off.astype(np.int64)
    assert recs == sorted((bytes(b[p[i]:p[i + 1]]), bytes(q[p[i]:p[i + 1]])) for i in range(50))
    k = int(h[0][1:])
    assert bytes(b[p[0]:p[1]]) == bytes(reads[0][o[k]:o[k + 1]])


def test_paired_mode_needs_two_inputs(capsys):
    assert parallel_cli.main(["only_one.fastq", "-p", "-t", "2", "-0"]) == 1


@pytest.mark.gpu
@pytest.mark.parametrize("paired", [False, True])
def test_cli_equals_oracle_blocks(tmp_path, paired):
    r1 = synth.make_reads(3000, 60, 21)
    r2 = synth.make_reads(3000, 60, 22)
    h1 = [b"@a%d" % i for i in range(3000)]
    f1, f2 = str(tmp_path / "m_1.fastq"), str(tmp_path / "m_2.fastq")
    fastq.write_fastq(f1, *r1, h1); fastq.write_fastq(f2, *r2)
    argv = [f1] + ([f2, "-p"] if paired else []) + ["-t", "4", "-0", "-H", "-M", "3", "-o", str(tmp_path / "out")]
    assert parallel_cli.main(argv) == 0

    def oracle_block(reads):
        b, q, l = H.oracle_build_ebwt(reads[0], reads[1], reads[2], 0)
        ob, oq, off, st, _, _, err = H.oracle_smooth_rebuild(b, q, l, H.make_params(M=3, B=0, eos=0))
        assert err == 0
        return ob, oq, off

    exp = parallel.run_blocks(r1, 4, oracle_block, r2 if paired else None)
    if paired:
        got1 = fastq.read_fastq(str(tmp_path / "out_1.fastq")); got2 = fastq.read_fastq(str(tmp_path / "out_2.fastq"))
        assert np.array_equal(got1[0], exp[0][0]) and np.array_equal(got1[1], exp[0][1]) and got1[3] == h1
        assert np.array_equal(got2[0], exp[1][0]) and np.array_equal(got2[1], exp[1][1])
    else:
        got = fastq.read_fastq(str(tmp_path / "out.fastq"))
        assert np.array_equal(got[0], exp[0]) and np.array_equal(got[1], exp[1]) and got[3] == h1
