#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv).
usage: python scripts/ncu_launches.py launches.csv"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, mi, ui, vi = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Unit"), H.index("Metric Value")
agg = collections.defaultdict(lambda: collections.defaultdict(float))
cnt = collections.Counter()
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    k = r[ki].split("(")[0]
    agg[k][r[mi]] += float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    if r[mi] == "gpu__time_duration.sum":
        cnt[k] += 1
tot = sum(a["gpu__time_duration.sum"] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1]["gpu__time_duration.sum"]):
    print(f"{k[-36:]:36s} launches {cnt[k]:4d}  time {a['gpu__time_duration.sum'] / 1e3:8.3f} ms  share {a['gpu__time_duration.sum'] / tot:5.3f}"
          f"  dram read {a['dram__bytes_read.sum'] / 1e6:8.1f} MB  write {a['dram__bytes_write.sum'] / 1e6:8.1f} MB")
print(f"total: {tot / 1e3:.3f} ms serialised, DRAM read {sum(a['dram__bytes_read.sum'] for a in agg.values()) / 1e9:.3f} GB, "
      f"write {sum(a['dram__bytes_write.sum'] for a in agg.values()) / 1e9:.3f} GB")
