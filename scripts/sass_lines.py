#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares of one kernel: joins the SASS page of an ncu report
(ncu -i X.ncu-rep --page source --csv) with the line table of the same build (nvdisasm -g -c of the cubin).

    python scripts/sass_lines.py <source.csv> <nvdisasm.txt> <mangled kernel prefix> <source file> [top]
"""
import collections
import csv
import re
import sys

csv_path, dis_path, prefix, src_path = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
lines = open(dis_path).read().split("\n")
start = [i for i, l in enumerate(lines) if l.startswith(".text." + prefix)][0]
amap, cur, curfile = [], None, None
for l in lines[start + 1:]:
    if l.startswith("\t.section") or l.startswith("//-----"):
        break
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        curfile, cur = m.group(1).split("/")[-1], int(m.group(2))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        amap.append((int(m.group(1), 16), curfile, cur))
rows = list(csv.reader(open(csv_path)))
h = rows[1]
ia, ii, isamp = h.index("Address"), h.index("Instructions Executed"), h.index("# Samples")
data = [r for r in rows[2:] if len(r) > ii]
base = int(data[0][ia], 16)
src = open(src_path).read().split("\n")
name = src_path.split("/")[-1]
inst, smp = collections.Counter(), collections.Counter()
for r, (off, f, ln) in zip(data, amap):
    assert int(r[ia], 16) - base == off
    inst[(f, ln)] += int(r[ii]); smp[(f, ln)] += int(r[isamp])
ti, ts = sum(inst.values()), sum(smp.values())
print("instructions", ti, "samples", ts)
for (f, ln), n in inst.most_common(top):
    text = src[ln - 1].strip()[:100] if f == name else ""
    print("%-24s %4d inst %4.1f%% smp %4.1f%%  %s" % (f, ln, 100 * n / ti, 100 * smp[(f, ln)] / ts, text))
if len(sys.argv) > 6:       # ranges "name:lo-hi,..." of the source file
    for spec in sys.argv[6].split(","):
        nm, rg = spec.split(":"); lo, hi = map(int, rg.split("-"))
        a = sum(n for (f, ln), n in inst.items() if f == name and lo <= ln <= hi)
        b = sum(n for (f, ln), n in smp.items() if f == name and lo <= ln <= hi)
        print("%-20s inst %5.1f%% smp %5.1f%%" % (nm, 100 * a / ti, 100 * b / ts))
