"""Summaries of an .ncu-rep read here with `ncu -i`: per-kernel pipe utilisation + stall reasons per issue, and
(with a kernel regex) the instructions that collect the most stall samples.
usage: python tools/ncu_summary.py file.ncu-rep [kernel-regex [cells-per-launch]]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
body = rows[2:]
names = [re.sub(r"\(.*", "", r[hdr.index("Kernel Name")]).replace("void ", "")[:34] for r in body]
print("kernels:", names)
keys = ["gpu__time_duration.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__warps_active.avg.per_cycle_active"]
for k in keys:
    if k in hdr:
        print("%-70s" % k, ["%.4g" % float(r[hdr.index(k)]) for r in body])
for h in hdr:
    m = re.match(r"smsp__average_warps_issue_stalled_(.*)_per_issue_active\.ratio$", h)
    if m:
        vals = [float(r[hdr.index(h)]) for r in body]
        if max(vals) > 0.15:
            print("stall %-22s" % m.group(1), ["%.2f" % v for v in vals])
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + sys.argv[2]],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hdr = next(r for r in rows if "Source" in r and "Instructions Executed" in r)
    i_s, i_e, i_w, i_a = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Address")
    seen, data = set(), []
    for r in rows:
        if r is hdr or len(r) <= i_e or not r[i_e].isdigit() or r[i_a] in seen:
            continue
        seen.add(r[i_a])
        data.append(r)
    cells = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    ops, st, tot = collections.Counter(), collections.Counter(), 0
    for d in data:
        toks = d[i_s].split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        op = ".".join(op.split(".")[:2]) if op.startswith(("MUFU", "UTC", "SYNCS", "LDT", "STS", "LDS", "STG", "LDG", "F2FP")) else op.split(".")[0]
        n = int(d[i_e])
        ops[op] += n
        tot += n
        st[op] += int(d[i_w] or 0)
    ts = sum(st.values())
    print("total warp instructions %d, per cell %.1f" % (tot, tot / cells))
    for op, n in ops.most_common(28):
        print("  %-14s %7.2f/cell  stall %5.1f%%" % (op, n / cells, 100.0 * st[op] / ts))
    print("top stall lines:")
    for r in sorted(data, key=lambda r: -int(r[i_w] or 0))[:12]:
        print("  %5.1f%% x%-9s %s" % (100.0 * int(r[i_w]) / ts, r[i_e], r[i_s][:70]))
