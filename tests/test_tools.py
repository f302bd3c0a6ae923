"""The two drop-in command-line tools (tools/bin), run with exactly the argv strings the reference
drivers build: BFQzip_ext.py:165-183 (step 1, eGap spelling), BFQzip.py:184 (gsufsort spelling)
and BFQzip_ext.py:207-215 (step 3, bfq_ext).  File contracts: .bwt / .bwt.qs / .1.lcp / .fq."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

import helpers as H
from bfqzip_b200 import fastq, synth

ROOT = H.ROOT
BIN = os.path.join(ROOT, "tools", "bin")


def _need_tools():
    if not os.path.exists(os.path.join(BIN, "eGap")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "tools"), "-j8"], stdout=subprocess.DEVNULL)


def _example(tmp_path):
    g = np.load(os.path.join(H.GOLDEN, "example.npz"))
    headers = g["headers"].tobytes().split(b"\n")
    fq = tmp_path / "reads.fastq"
    fastq.write_fastq(str(fq), g["bases"], g["quals"], g["off"], headers)
    return g, headers, fq


def test_tools_fail_loudly_without_gpu(tmp_path):
    """No CPU fallback: without a device both tools exit non-zero, which the driver maps to failure."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _need_tools()
    g, headers, fq = _example(tmp_path)
    r = subprocess.run([os.path.join(BIN, "eGap"), str(fq), "--em", "--mem", "0", "--qs", "-o", str(tmp_path / "o"), "--lcp", "--lbytes", "1"],
                       capture_output=True)
    assert r.returncode != 0 and b"no CPU path" in r.stderr
    for e, a in ((".bwt", g["bwt"]), (".bwt.qs", g["qs"]), (".1.lcp", g["lcp"])):
        a.tofile(str(tmp_path / ("o" + e)))
    r = subprocess.run([os.path.join(BIN, "bfq_ext"), "-e", "o.bwt", "-q", "o.bwt.qs", "-a", "o.1.lcp", "-o", "o", "-l", "250", "-s", "35", "-m", "5"],
                       cwd=tmp_path, capture_output=True)
    assert r.returncode != 0 and not os.path.exists(tmp_path / "o.fq")


def test_bfq_ext_help_exit_code():
    """The reference prints its help and exits 0 when a required flag is missing (bfq_ext.cpp:104-129, 1033)."""
    _need_tools()
    r = subprocess.run([os.path.join(BIN, "bfq_ext"), "-e", "x", "-q", "y"], capture_output=True)
    assert r.returncode == 0 and b"bfq_ext [options]" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("style", ["eGap", "gsufsort"])
def test_driver_command_lines_example(tmp_path, style):
    """Config #1: example/reads.fastq through step 1 + step 3 as the driver spells them (--m3 => headers)."""
    _need_tools()
    g, headers, fq = _example(tmp_path)
    out = str(tmp_path / "out")
    if style == "eGap":
        cmd = f"{BIN}/eGap {fq} --em --mem 0 --qs -o {out} --lcp --lbytes 1"
        eos = 0
    else:
        cmd = f"{BIN}/gsufsort {fq} --bwt --qs --lcp 1 -o {out}"
        eos = 35
    subprocess.check_call(cmd.split(), stdout=subprocess.DEVNULL)
    rd = lambda e: np.fromfile(out + e, dtype=np.uint8)
    is_eos = g["bwt"] == 35
    eb = g["bwt"].copy(); eb[is_eos] = eos
    eq = g["qs"].copy(); eq[is_eos] = eos
    assert np.array_equal(rd(".bwt"), eb) and np.array_equal(rd(".bwt.qs"), eq) and np.array_equal(rd(".1.lcp"), g["lcp"])
    with open(out + ".h", "wb") as f:                      # step 2: sed -n 1~4p (BFQzip_ext.py:185-196)
        f.write(b"".join(h + b"\n" for h in headers))
    cmd = f"{BIN}/bfq_ext -e {out}.bwt -q {out}.bwt.qs -a {out}.1.lcp -o {out} -l 250 -s {eos} -m 5 -H {out}.h"
    r = subprocess.run(cmd.split(), cwd=tmp_path, capture_output=True, check=True)
    data = open(out + ".fq", "rb").read()
    assert hashlib.md5(data).hexdigest() == "9178301c8ef6c9864d5ebfccf47313d4"
    st = H.parse_ref_stats(r.stdout.decode())
    assert (st["clusters"], st["qs_smoothed"], st["modified"]) == (387, 4198, 10)
    # checkFASTQ.py:18-32: every record has len(dna) == len(qs)
    lines = data.split(b"\n")
    assert all(len(lines[i + 1]) == len(lines[i + 3]) for i in range(0, len(lines) - 1, 4))
    assert not os.path.exists(tmp_path / "tmp")            # no ./tmp scratch needed


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [(0, 0), (1, 1), (3, 0), (2, 1)], ids=lambda m: f"M{m[0]}B{m[1]}")
def test_bfq_ext_modes_vs_oracle(tmp_path, mode):
    M, B = mode
    _need_tools()
    bases, quals, off = synth.make_reads(3000, 80, 77 + M, var_len=True)
    fq = tmp_path / "in.fastq"
    fastq.write_fastq(str(fq), bases, quals, off)
    out = str(tmp_path / "o")
    subprocess.check_call(f"{BIN}/eGap {fq} --em --mem 1 --qs -o {out} --lcp --lbytes 1".split(), stdout=subprocess.DEVNULL)
    r = subprocess.run(f"{BIN}/bfq_ext_M{M}_B{B} -e {out}.bwt -q {out}.bwt.qs -a {out}.1.lcp -o {out} -l 250 -s 0 -m 5 -k 12 -v 70".split(),
                       capture_output=True, check=True)
    rb, rq, rl = H.oracle_build_ebwt(bases, quals, off, 0)
    b, q, o, st, _, _, err = H.oracle_smooth_rebuild(rb, rq, rl, H.make_params(M=M, B=B, K=12, eos=0, default_q=70))
    assert err == 0
    assert open(out + ".fq", "rb").read() == fastq.format_fastq(b, q, o)
    got = H.parse_ref_stats(r.stdout.decode())
    assert all(got[k] == st[k] for k in got)


@pytest.mark.gpu
def test_bfq_ext_component_streams_and_host_paths(tmp_path):
    """-C writes OUT.fq.dna / OUT.fq.qs = `sed -n 2~4p` / `sed -n 4~4p` of OUT.fq (step 4's streams, BFQzip_ext.py:19-21);
    the device-side FASTQ parse / emit of the tools and their host-side fallbacks give identical files."""
    _need_tools()
    bases, quals, off = synth.make_reads(4000, 70, 5, var_len=True)
    fq = tmp_path / "in.fastq"
    fastq.write_fastq(str(fq), bases, quals, off)
    outs = {}
    for tag, env in (("dev", {}), ("host", {"BFQ_HOST_PARSE": "1", "BFQ_HOST_EMIT": "1"})):
        out = str(tmp_path / tag)
        e = dict(os.environ, **env)
        subprocess.check_call(f"{BIN}/eGap {fq} --em --mem 1 --qs -o {out} --lcp --lbytes 1".split(), stdout=subprocess.DEVNULL, env=e)
        subprocess.check_call(f"{BIN}/bfq_ext -e {out}.bwt -q {out}.bwt.qs -a {out}.1.lcp -o {out} -l 250 -s 0 -m 5".split(),
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, env=e)
        outs[tag] = [open(out + ext, "rb").read() for ext in (".bwt", ".bwt.qs", ".1.lcp", ".fq")]
    assert outs["dev"] == outs["host"]
    out = str(tmp_path / "dev")
    subprocess.check_call(f"{BIN}/bfq_ext -e {out}.bwt -q {out}.bwt.qs -a {out}.1.lcp -o {out} -l 250 -s 0 -m 5 -C".split(),
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    lines = open(out + ".fq", "rb").read().split(b"\n")
    assert open(out + ".fq", "rb").read() == outs["dev"][3]
    assert open(out + ".fq.dna", "rb").read() == b"".join(x + b"\n" for x in lines[1::4])
    assert open(out + ".fq.qs", "rb").read() == b"".join(x + b"\n" for x in lines[3::4])
