"""per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list
    python profiles/tools/launch_summary.py launches.csv [last_n]"""
import csv
import re
import sys

with open(sys.argv[1]) as f:
    rows = list(csv.DictReader(l for l in f if not l.startswith("==")))
seq = [(re.sub(r"\(.*", "", r["Kernel Name"])[:64], float(r["Metric Value"].replace(",", "")))
       for r in rows]
if len(sys.argv) > 2:
    seq = seq[-int(sys.argv[2]):]
agg = {}
for n, t in seq:
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += t
tot = sum(v[1] for v in agg.values())
print(f"{len(seq)} launches, {tot / 1e3:.1f} us (cold-cache, serialised: compare shares)")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{n:64s} {c:4d} {t / 1e3:9.1f} us {t / c / 1e3:8.1f} us/launch {100 * t / tot:5.1f}%")
