#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total, share, avg.
    python tools/summarize_launches.py gpurun_out/launches.csv [--skip N] [--md]"""
import csv, sys, re, collections
path = sys.argv[1]
skip = int(sys.argv[sys.argv.index("--skip") + 1]) if "--skip" in sys.argv else 0
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if not l.startswith("==")]
rd = csv.DictReader(lines)
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    us = v / 1e3 if unit in ("ns", "nsecond") else v if unit in ("us", "usecond") else v * 1e3
    name = re.sub(r"\(.*$", "", r["Kernel Name"])
    rows.append((name, us, r.get("Grid Size", ""), r.get("Block Size", "")))
rows = rows[skip:]
agg = collections.OrderedDict()
for n, us, *_ in rows:
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values())
print(f"launches {len(rows)}  total {tot:.1f} us")
print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"| `{n[:90]}` | {c} | {t:.1f} | {100*t/tot:.1f}% | {t/c:.2f} |")
