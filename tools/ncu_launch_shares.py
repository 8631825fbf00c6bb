"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel:
python tools/ncu_launch_shares.py <launches.csv>"""
import collections
import csv
import sys

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
agg = collections.OrderedDict()
for r in rows:
    a = agg.setdefault(r['Kernel Name'][:90], [0, 0.0])
    a[0] += 1; a[1] += float(r['Metric Value']) / 1e3
tot = sum(v[1] for v in agg.values())
print(f"{len(rows)} launches, {tot:.1f} us total (cold-cache, serialised: compare shares)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1]:10.1f} us {100 * v[1] / tot:5.1f}%  x{v[0]:4d}  avg {v[1] / v[0]:8.1f} us  {k}")
