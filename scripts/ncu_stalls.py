#!/usr/bin/env python
"""Top stalled SASS instructions of one kernel launch in an ncu report, with the dominant stall reasons.
usage: python scripts/ncu_stalls.py file.ncu-rep kernel_regex [top_n]"""
import csv, io, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[1]
si, sm, ei = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
names = ["stall_long_sb", "stall_wait", "stall_short_sb", "stall_math", "stall_lg", "stall_mio", "stall_branch_resolving",
         "stall_not_selected", "stall_no_inst", "stall_dispatch", "stall_membar", "stall_sleep", "stall_selected"]
idx = {n: h.index(n) for n in names if n in h}
data = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name" and data:
        break                      # first launch only
    if len(r) == len(h) and r[sm].isdigit():
        data.append(r)
g = lambda r, i: int(r[i] or 0)
tot = sum(g(r, sm) for r in data)
print("kernel", rows[0][1][:80], "| total samples", tot, "| warp-inst", sum(g(r, ei) for r in data))
print({n[6:]: round(100 * sum(g(r, i) for r in data) / max(1, tot), 1) for n, i in idx.items()})
for r in sorted(data, key=lambda r: -g(r, sm))[:top_n]:
    why = " ".join(n[6:12] + ":" + r[i] for n, i in idx.items() if g(r, i) > 0.15 * g(r, sm))
    print(r[sm].rjust(6), r[ei].rjust(9), why.ljust(40), "|", r[si].strip()[:90])
