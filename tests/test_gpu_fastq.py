"""FASTQ text parsed and written on the device (bfq_*_fastq entry points) against the host-side reference semantics
(bfqzip_b200/fastq.py mirrors the loaders and the writer of the reference) and against the flat-array entry points."""
import numpy as np
import pytest

from bfqzip_b200 import fastq, lib, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g():
    x = lib.BfqGpu(2, 0, 0)
    yield x
    x.close()


def host_reads(text):
    b, q, off, _ = fastq.parse_fastq_bytes(text)
    keep = np.diff(off.astype(np.int64)) > 0                 # cat_all drops empty reads (lib/utils.c:229-236)
    if not keep.all():
        o = off.astype(np.int64)
        b = np.concatenate([b[o[i]:o[i + 1]] for i in range(len(keep)) if keep[i]] or [np.empty(0, np.uint8)])
        q = np.concatenate([q[o[i]:o[i + 1]] for i in range(len(keep)) if keep[i]] or [np.empty(0, np.uint8)])
        off = np.concatenate([[0], np.cumsum(np.diff(o)[keep])]).astype(np.uint64)
    return b, q, off


def check_text(g, text, eos=0):
    b, q, off = host_reads(text)
    want = g.build_ebwt(b, q, off, eos)
    bwt, qs, lcp, nreads = g.build_ebwt_fastq(text, eos)
    assert nreads == len(off) - 1
    assert np.array_equal(bwt, want[0]) and np.array_equal(qs, want[1]) and np.array_equal(lcp, want[2])
    return b, q, off, (bwt, qs, lcp)


def test_ingest_matches_host_parser(g):
    b, q, off = synth.make_reads(30000, 100, 7)
    check_text(g, fastq.format_fastq(b, q, off))


def test_ingest_variable_lengths_headers_and_last_line_without_newline(g):
    b, q, off = synth.make_reads(5000, 90, 8, var_len=True)
    heads = [b"@read%d some text + more @ signs" % i for i in range(len(off) - 1)]
    text = fastq.format_fastq(b, q, off, heads)
    check_text(g, text)
    check_text(g, text[:-1])                    # no trailing newline
    check_text(g, text + b"@partial\nACGT\n")   # trailing partial record is ignored (lib/file.c:182-227)


def test_ingest_crlf_and_empty_reads(g):
    recs = [(b"ACGTAC", b"IIIIII"), (b"", b""), (b"GGGTTTAAC", b"#########"), (b"T", b"!")]
    text = b"".join(b"@r\r\n" + s + b"\r\n+\r\n" + qq + b"\r\n" for s, qq in recs)
    b, q, off, _ = check_text(g, text, eos=35)
    assert len(off) - 1 == 3 and bytes(b) == b"ACGTACGGGTTTAACT"


def test_ingest_rejects_mismatched_lengths(g):
    with pytest.raises(lib.BfqError):
        g.build_ebwt_fastq(b"@a\nACGT\n+\nIII\n", 0)


def test_ingest_empty_and_tiny(g):
    for text in (b"", b"\n", b"@a\nA\n+\n"):
        bwt, qs, lcp, nreads = g.build_ebwt_fastq(text, 0)
        assert nreads == 0 and len(bwt) == 0
    bwt, qs, lcp, nreads = g.build_ebwt_fastq(b"@a\nA\n+\nI", 0)
    assert nreads == 1 and bytes(bwt) == b"A\x00"


def test_emit_matches_host_writer(g):
    b, q, off = synth.make_reads(8000, 80, 9, var_len=True)
    P = lib.default_params(eos=0)
    bwt, qs, lcp = g.build_ebwt(b, q, off, 0)
    ob, oq, ooff, st = g.smooth_rebuild(bwt, qs, lcp, P)
    text, st2 = g.smooth_rebuild_fastq(bwt, qs, lcp, P)
    assert text == fastq.format_fastq(ob, oq, ooff) and st2 == st
    # the component streams of step 4: sed -n 2~4p / sed -n 4~4p of the FASTQ just written
    text3, _, dna, qsl = g.smooth_rebuild_fastq(bwt, qs, lcp, P, components=True)
    lines = text.split(b"\n")
    assert text3 == text
    assert dna == b"".join(x + b"\n" for x in lines[1::4]) and qsl == b"".join(x + b"\n" for x in lines[3::4])


def test_pipeline_fastq_text_to_text(g):
    b, q, off = synth.make_reads(40000, 100, 10)
    P = lib.default_params(eos=0)
    (bwt, qs, lcp), (ob, oq), st = g.pipeline(b, q, off, P)
    files, text, st2 = g.pipeline_fastq(fastq.format_fastq(b, q, off), P)
    assert st2 == st and text == fastq.format_fastq(ob, oq, off)
    assert np.array_equal(files[0], bwt) and np.array_equal(files[1], qs) and np.array_equal(files[2], lcp)
    _, text3, _ = g.pipeline_fastq(fastq.format_fastq(b, q, off), P, want_files=False)
    assert text3 == text
