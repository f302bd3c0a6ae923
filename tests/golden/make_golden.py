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


# ---------------------------------------------------------------------------------------------------------------
# --large: BASELINE-size cases.  The arrays are too large to commit, so tests/golden/large.json holds md5 digests of
# what the UNMODIFIED reference binaries write (gsufsort main.c:848-895: .bwt/.bwt.qs/.1.lcp; eGap for the EOS-0
# spelling; bfq_ext decode.cpp:1040-1127: the .fq, split here into its base and quality lines) plus the seven
# counters of bfq_ext.cpp:1113-1128.  Inputs are regenerated from the seeds by tests/large_cases.py on the GPU box.
def run_large(only=None):
    import hashlib, json, shutil, subprocess, time
    import large_cases as LC
    md5 = lambda a: hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()
    path = os.path.join(HERE, "large.json")
    res = json.load(open(path)) if os.path.exists(path) else {}
    threads = str(os.cpu_count() or 1)
    for name, case in LC.CASES.items():
        if only and name not in only:
            continue
        t0 = time.time()
        bases, quals, off = LC.make_case(name)
        M, B = case["M"], case["B"]
        wd = tempfile.mkdtemp(prefix="bfqlarge_", dir=os.environ.get("BFQ_TMP", "/tmp"))
        fq = os.path.join(wd, "in.fastq")
        with open(fq, "wb") as f:
            f.write(LC.format_fastq_fast(bases, quals, off))
        n = int(off[-1]) + len(off) - 1
        exe = "gsufsort-64" if n > (1 << 30) else "gsufsort"
        out = os.path.join(wd, "out")
        subprocess.check_call([os.path.join(H.REF_DIR, exe), fq, "--bwt", "--qs", "--lcp", "1", "-o", out],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        t1 = time.time()
        rd = lambda e: np.fromfile(out + e, dtype=np.uint8)
        bwt, qs, lcp = rd(".bwt"), rd(".bwt.qs"), rd(".1.lcp")
        r = dict(gen=case, nreads=len(off) - 1, n=n, in_bases=md5(bases), in_quals=md5(quals),
                 eos35=dict(bwt=md5(bwt), qs=md5(qs), lcp=md5(lcp)))
        eos = 35
        if case.get("egap"):
            # the driver's own step 1 (BFQzip_ext.py:173-177): eGap --em --mem MB --qs --lcp --lbytes 1, EOS byte 0
            eg = os.path.join(wd, "eg")
            subprocess.check_call([sys.executable, os.path.join(H.REF_DIR, "egap", "eGap"), fq, "--em", "--mem",
                                   str(max(1, os.path.getsize(fq) >> 20)), "--qs", "-o", eg, "--lcp", "--lbytes", "1"],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=wd)
            rde = lambda e: np.fromfile(eg + e, dtype=np.uint8)
            bwt, qs, lcp = rde(".bwt"), rde(".bwt.qs"), rde(".1.lcp")
            r["eos0"] = dict(bwt=md5(bwt), qs=md5(qs), lcp=md5(lcp))
            eos = 0
        t2 = time.time()
        del bases, quals
        for a, nm in ((bwt, "x.bwt"), (qs, "x.bwt.qs"), (lcp, "x.1.lcp")):
            a.tofile(os.path.join(wd, nm))
        del bwt, qs, lcp
        rr = subprocess.run([os.path.join(H.REF_DIR, f"bfq_ext_M{M}_B{B}"), "-e", "x.bwt", "-q", "x.bwt.qs", "-a", "x.1.lcp",
                             "-o", "x", "-l", "250", "-s", str(eos), "-m", "5", "-T", threads], cwd=wd,
                            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
        assert rr.returncode == 0, rr.returncode
        st = H.parse_ref_stats(rr.stdout.decode("latin1"))
        with open(os.path.join(wd, "x.fq"), "rb") as f:
            text = np.frombuffer(f.read(), np.uint8)
        ob, oq = LC.parse_fastq_fast(text, off)
        r[f"M{M}_B{B}"] = dict(eos=eos, fq=md5(text), out_bases=md5(ob), out_quals=md5(oq),
                              stats={k: st[k] for k in STAT_KEYS})
        r["cpu_seconds"] = dict(step1=round(t1 - t0, 1), egap=round(t2 - t1, 1), bfq_ext=round(time.time() - t2, 1),
                                threads=int(threads))
        shutil.rmtree(wd, ignore_errors=True)
        res[name] = r
        with open(path, "w") as f:
            json.dump(res, f, indent=1, sort_keys=True)
        print(name, json.dumps(r["cpu_seconds"]), flush=True)


if __name__ == "__main__":
    if "--large" in sys.argv:
        run_large([a for a in sys.argv[1:] if not a.startswith("--")])
        sys.exit(0)
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
