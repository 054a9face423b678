"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count, total time, share."""
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith("==")) if len(r) > 14]
hdr, rows = rows[0], rows[1:]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = defaultdict(lambda: [0, 0.0, 0.0])
for r in rows:
    name = r[ki].split("(")[0]
    ns = float(r[vi].replace(",", ""))
    a = agg[name]
    a[0] += 1; a[1] += ns; a[2] = max(a[2], ns)
tot = sum(a[1] for a in agg.values())
print(f"| kernel | launches | total ms | share | max launch ms |\n|---|---|---|---|---|")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k}` | {a[0]} | {a[1] / 1e6:.3f} | {100 * a[1] / tot:.2f}% | {a[2] / 1e6:.3f} |")
print(f"\ntotal {tot / 1e6:.1f} ms over {len(rows)} launches")
