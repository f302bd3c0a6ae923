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


DRIVER = os.path.join(H.REF_DIR, "driver")


def _driver_tree(tmp_path):
    """The tree INTEGRATION.md section A describes: the reference's UNMODIFIED BFQzip_ext.py + checkFASTQ.py (copied into
    git-ignored oracle/_ref/driver/ by oracle/Makefile) with the two drop-in tools where the driver looks for its own
    (it resolves them relative to its directory, BFQzip_ext.py:55,173,207)."""
    d = tmp_path / "bfqzip"
    (d / "external" / "egap").mkdir(parents=True)
    (d / "src_ext_mem").mkdir()
    for f in ("BFQzip_ext.py", "checkFASTQ.py"):
        (d / f).write_bytes(open(os.path.join(DRIVER, f), "rb").read())
    os.symlink(os.path.join(BIN, "eGap"), d / "external" / "egap" / "eGap")
    os.symlink(os.path.join(BIN, "bfq_ext"), d / "src_ext_mem" / "bfq_ext")
    # Step 5 (BFQzip_ext.py:248-272) hands the streams to the host entropy coders `7z a -mm=PPMd` and external/libbsc/bsc
    # (resolved relative to the cwd).  Both are out of scope and neither exists on the GPU box; a missing 7z
    # would end the driver with an uncaught FileNotFoundError before its final report.  Stand-ins that COPY their input let
    # the unmodified driver run to its end; nothing of steps 1-4 depends on them.
    stub = tmp_path / "stubs"
    (stub).mkdir()
    (stub / "7z").write_text("#!/bin/sh\n# 7z a -mm=PPMd OUT IN\ncp \"$4\" \"$3\"\n")
    (d / "external" / "libbsc").mkdir()
    (d / "external" / "libbsc" / "bsc").write_text("#!/bin/sh\n# bsc e IN OUT -T\ncp \"$2\" \"$3\"\n")
    os.chmod(stub / "7z", 0o755); os.chmod(d / "external" / "libbsc" / "bsc", 0o755)
    return d, stub


@pytest.mark.gpu
def test_unmodified_reference_driver_runs_end_to_end(tmp_path):
    """BASELINE config 1 as it is DEFINED: `python3 BFQzip_ext.py reads.fastq -o out --m3` with the reference's own driver
    (BFQzip_ext.py:165-244, 310-318): step 1 (our eGap, --mem 0 on this 24 KB input), step 2 (headers), step 3 (our
    bfq_ext with -H), step 4 (component streams), step 5 (stand-ins), final report; then checkFASTQ.py:18-32 on out.fq."""
    if not os.path.exists(os.path.join(DRIVER, "BFQzip_ext.py")):
        pytest.skip("oracle/_ref/driver is not built (make -C oracle ref, needs /root/reference)")
    _need_tools()
    import sys
    g, headers, _ = _example(tmp_path)
    d, stub = _driver_tree(tmp_path)
    fq = d / "reads.fastq"
    fastq.write_fastq(str(fq), g["bases"], g["quals"], g["off"], headers)
    env = dict(os.environ, PATH=str(stub) + os.pathsep + os.environ.get("PATH", ""))
    r = subprocess.run([sys.executable, "BFQzip_ext.py", "reads.fastq", "-o", "out", "--m3"], cwd=d, env=env, capture_output=True)
    log = (d / "out.log").read_text(errors="replace") if (d / "out.log").exists() else ""
    assert r.returncode == 0, (r.stdout.decode(errors="replace")[-2000:], r.stderr.decode(errors="replace")[-2000:], log[-2000:])
    stdout = r.stdout.decode()
    assert "Valid FASTQ file!" in stdout and "=== results ===" in stdout
    assert "--mem 0 " in stdout                                        # the driver asks eGap for 0 MB on small files
    # step 1: the three files, byte-identical to the reference's (EOS byte 0 in the eGap spelling)
    bwt = g["bwt"].copy(); bwt[g["bwt"] == 35] = 0
    qs = g["qs"].copy(); qs[g["bwt"] == 35] = 0
    assert np.array_equal(np.fromfile(d / "out.bwt", np.uint8), bwt)
    assert np.array_equal(np.fromfile(d / "out.bwt.qs", np.uint8), qs)
    assert np.array_equal(np.fromfile(d / "out.1.lcp", np.uint8), g["lcp"])
    # step 3: the .fq of the reference, its statistics block in the log
    assert hashlib.md5((d / "out.fq").read_bytes()).hexdigest() == "9178301c8ef6c9864d5ebfccf47313d4"
    st = H.parse_ref_stats(log)
    assert (st["clusters"], st["qs_smoothed"], st["modified"]) == (387, 4198, 10)
    # steps 2 and 4 ran on their inputs; step 5's stand-ins produced what the final report sizes up
    assert (d / "out.h").read_bytes() == b"".join(h + b"\n" for h in headers)
    for f in ("out.fq.dna", "out.fq.qs", "out.h.7z", "out.fq.dna.7z", "out.fq.qs.7z", "out.fq.qs.bsc"):
        assert (d / f).exists(), f
    # the reference's own validity check on the result
    sys.path.insert(0, str(d))
    try:
        import importlib
        checker = importlib.import_module("checkFASTQ")
        assert checker.checkFASTQ(str(d / "out.fq")) is True
    finally:
        sys.path.remove(str(d))
        sys.modules.pop("checkFASTQ", None)
