#!/usr/bin/env python
"""Hot SASS regions of an `ncu --page source --csv` dump: contiguous instruction ranges with similar
execution counts, with their share of executed instructions and of stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
ie = h.index('Instructions Executed'); ws = h.index('Warp Stall Sampling (All Samples)')
data = [(i, r[1].strip(), int(r[ie]), int(r[ws])) for i, r in enumerate(rows[2:]) if len(r) > ie]
tot = sum(d[2] for d in data); tw = sum(d[3] for d in data)
print("total inst", tot, "samples", tw, "sass lines", len(data))
reg = []
for d in data:
    if reg and abs(d[2] - reg[-1][2]) <= 0.3 * max(d[2], reg[-1][2], 1):
        reg[-1][1] = d[0]; reg[-1][3] += d[2]; reg[-1][4] += d[3]; reg[-1][2] = d[2]
    else:
        reg.append([d[0], d[0], d[2], d[2], d[3]])
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
for r in reg:
    if r[3] > thr * tot:
        ops = [data[i][1].split()[0] for i in range(r[0], min(r[1] + 1, r[0] + 400))]
        import collections
        top = collections.Counter(o.split('.')[0] for o in ops).most_common(6)
        print(f"idx {r[0]:4d}-{r[1]:4d} n={r[1]-r[0]+1:3d} exec~{r[2]/1e6:8.1f}M inst {100*r[3]/tot:5.1f}% stall {100*r[4]/tw:5.1f}%  {top}")
