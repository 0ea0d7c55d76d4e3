#!/usr/bin/env python
"""Executed warp-instructions / stall samples per CUDA source line for one kernel.
Joins the SASS page of an ncu report (instruction order) with `nvdisasm -g` line info of the built
library.  usage: python scripts/ncu_lines.py file.ncu-rep kernel_regex mangled_substr [top_n] [launch_idx]"""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, pat, mangled = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60
which = int(sys.argv[5]) if len(sys.argv) > 5 else 0
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "trajectorycrafter_b200", "libtcwarp.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
cubin = [os.path.join(tmp, f) for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
# walk the disassembly of the wanted function: remember current //## File ..., line N and list instructions
insts, cur, infn = [], ("?", 0), False
for ln in dis.splitlines():
    if ln.startswith("//---") and ".text." in ln:
        infn = mangled in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        insts.append((cur, m.group(2).strip()))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
blocks, curb = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        curb = {"name": row[1], "hdr": None, "rows": []}; blocks.append(curb)
    elif curb is not None and curb["hdr"] is None:
        curb["hdr"] = row
    elif curb is not None and row:
        curb["rows"].append(row)
b = blocks[which]; h = b["hdr"]
ei, sm = h.index("Instructions Executed"), h.index("# Samples")
rows = b["rows"]
print(b["name"][:100], "| sass rows", len(rows), "| nvdisasm insts", len(insts))
n = min(len(rows), len(insts))
per_line, per_line_s, ops = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
tot = stot = 0
for i in range(n):
    e, s = int(rows[i][ei] or 0), int(rows[i][sm] or 0)
    per_line[insts[i][0]] += e; per_line_s[insts[i][0]] += s; tot += e; stot += s
    t = insts[i][1].split(); op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    ops[insts[i][0]][op] += e
div = float(os.environ.get("PER", "1"))
print(f"total warp-inst {tot}  (divide by PER={div}: {tot/div:.1f})  samples {stot}")
for key, e in per_line.most_common(top):
    o = " ".join(f"{k}:{v/div:.1f}" for k, v in ops[key].most_common(6))
    print(f"{key[0]}:{key[1]:<4} {e/div:8.1f} {100*e/tot:5.1f}% smp {100*per_line_s[key]/max(1,stot):5.1f}%  {o}")
