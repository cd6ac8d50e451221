"""Instruction mix of one kernel from an `ncu --page source --csv --print-source sass` export.
usage: python tools/sass_mix.py export.csv <cells-per-launch>"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
cells = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr = next(r for r in rows if "Source" in r and "Instructions Executed" in r)
i_s, i_e, i_w = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
ops, stalls, tot = collections.Counter(), collections.Counter(), 0
for d in rows:
    if d is hdr or len(d) <= i_e or not d[i_e].isdigit():
        continue
    toks = d[i_s].split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    keep2 = op.startswith(("MUFU", "F2F", "UTC", "SYNCS", "LDT", "STS", "LDS", "STG", "LDG", "F2FP"))
    op = ".".join(op.split(".")[:2]) if keep2 else op.split(".")[0]
    n = int(d[i_e])
    ops[op] += n
    tot += n
    stalls[op] += int(d[i_w] or 0)
print("total warp instructions", tot, " per warp-cell %.1f" % (tot / cells))
ts = sum(stalls.values())
for op, n in ops.most_common(45):
    print("%-16s %11d  %6.2f/cell  stall samples %5.1f%%" % (op, n, n / cells, 100.0 * stalls[op] / max(ts, 1)))
