import sys, os, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import helpers as H
from bfqzip_b200 import lib
name = sys.argv[1] if len(sys.argv) > 1 else "example"
M = int(sys.argv[2]) if len(sys.argv) > 2 else 0
g = np.load(f"tests/golden/{name}.npz")
c = lib.BfqGpu(M, 0, 0)
ob, oq, off, st = c.smooth_rebuild(g["bwt"], g["qs"], g["lcp"], lib.default_params(eos=35))
q, b = c.export_rows(len(g["bwt"]))
sq, obw, est, err = H.oracle_smooth(g["bwt"], g["qs"], g["lcp"], H.make_params(M=M, B=0))
rows = g["bwt"] != 35
bad = np.nonzero((q != sq) & rows)[0]
print("stats gpu", st); print("stats ora", est)
print("row mismatches", len(bad), "base mismatches", int((b != obw).sum()))
lcp = g["lcp"]
for i in bad[:12]:
    lo, hi = max(0, i - 6), min(len(lcp), i + 4)
    print("row", i, "exp", sq[i], "got", q[i], "orig", g["qs"][i])
    print("   lcp", lcp[lo:hi].tolist()); print("   bwt", bytes(g["bwt"][lo:hi])); print("   qs ", g["qs"][lo:hi].tolist()); print("   exp", sq[lo:hi].tolist()); print("   got", q[lo:hi].tolist())
print("final quals equal:", np.array_equal(oq, g[f"out_quals_M{M}_B0"]))
