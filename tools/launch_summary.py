#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv).
    python tools/launch_summary.py launches.csv "title line" > summary.txt"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]; ki = hdr.index("Kernel Name"); vi = hdr.index("Metric Value"); ui = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    v = v / 1000.0 if r[ui] == "ns" else v * 1000.0 if r[ui] == "ms" else v
    a = agg.setdefault(r[ki].split("(")[0][:48], [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(a[1] for a in agg.values())
print((sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]) + "\n")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:50s} n={a[0]:5d} total {a[1]:10.1f} us  avg {a[1] / a[0]:9.1f} us  {100 * a[1] / tot:5.1f}%")
