#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source cuda,sass` of one kernel: stall samples and executed
warp-instructions by CUDA source line and by SASS opcode.  usage: ncu_sass_hot.py report.ncu-rep kernel_regex [top] [launch#]"""
import collections
import csv
import io
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + rx],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
by_line, by_op, src = collections.Counter(), collections.Counter(), {}
by_line_e, by_op_e = collections.Counter(), collections.Counter()
fname, H, seen_funcs = "", None, 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        seen_funcs += 1
        continue
    if r[0] == "Line No":
        H = r
        li, ci, ai = 0, 1, 2
        si = 3
        ni, ei = H.index("# Samples"), H.index("Instructions Executed")
        continue
    if H is None or len(r) <= ei:
        continue
    if r[li]:
        cur = (fname, int(r[li]))
        src[cur] = r[ci].strip()
    if not r[ai]:
        continue
    try:
        s, e = int(r[ni]), int(r[ei])
    except ValueError:
        continue
    ins = r[si].strip()
    parts = ins.split()
    op = (parts[1] if parts[0].startswith("@") and len(parts) > 1 else parts[0]).split(".")[0]
    by_line[cur] += s
    by_line_e[cur] += e
    by_op[op] += s
    by_op_e[op] += e
ts, te = sum(by_op.values()), sum(by_op_e.values())
print("%s: %d samples, %d warp-instructions (all captured launches of the kernel)" % (rx, ts, te))
print("by opcode:")
for op, s in by_op.most_common(16):
    print("  %-10s samples %7d (%4.1f%%)  executed %11d (%4.1f%%)" % (op, s, 100.0 * s / ts, by_op_e[op], 100.0 * by_op_e[op] / te))
print("by source line:")
for k, s in by_line.most_common(top):
    print("  %-16s:%4d samples %7d (%4.1f%%) exec %10d (%4.1f%%) | %s" % (k[0], k[1], s, 100.0 * s / ts, by_line_e[k], 100.0 * by_line_e[k] / te, src.get(k, "")[:100]))
