#!/usr/bin/env python
"""Per-opcode executed-instruction totals from an ncu report's source page (SASS view).
usage: python scripts/ncu_source.py file.ncu-rep [launch_index]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None and row:
        cur["rows"].append(row)
b = blocks[which]
h = b["hdr"]
si, ei = h.index("Source"), h.index("Instructions Executed")
sm = h.index("# Samples")
warps = max(int(r[ei]) for r in b["rows"])
print(b["name"][:90], "| launches in file:", len(blocks), "| max exec count (=warps):", warps)
ops = collections.Counter(); tot = 0; samples = collections.Counter()
for r in b["rows"]:
    toks = r[si].split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    op = op.split(".")[0].rstrip(";")
    n = int(r[ei]); ops[op] += n; tot += n; samples[op] += int(r[sm])
stot = sum(samples.values())
print(f"total warp-instructions {tot}  = {tot / warps:.1f} per warp")
for op, n in ops.most_common(30):
    print(f"  {op:12s} {n / warps:7.1f} /warp   {100 * n / tot:5.1f}%   samples {100 * samples[op] / max(1, stot):5.1f}%")
