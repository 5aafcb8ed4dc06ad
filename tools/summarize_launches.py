"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table (markdown)."""
import csv
import sys
from collections import OrderedDict

path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    val = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = val / 1000.0 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1000.0)
    rows.append((r["Kernel Name"], r.get("Grid Size", ""), r.get("Block Size", ""), us))
agg = OrderedDict()
for name, grid, block, us in rows:
    short = name.split("(")[0].replace("void ", "").replace("rtdc::", "")
    key = (short, grid, block)
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += us
total = sum(v[1] for v in agg.values())
print(f"| kernel | grid | block | launches | total us | mean us | share |")
print("|---|---|---|---:|---:|---:|---:|")
for (name, grid, block), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{name[:70]}` | {grid} | {block} | {n} | {us:.1f} | {us / n:.1f} | {100 * us / total:.1f}% |")
print(f"\ntotal {total:.1f} us over {len(rows)} launches (serialised under ncu: compare shares, not absolutes)")
