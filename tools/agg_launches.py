"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (last N launches = one step)."""
import collections
import csv
import re
import sys

path = sys.argv[1]
last = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rows = list(csv.reader(open(path, errors="ignore")))
hdr = None
items = []
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get("Metric Name") == "gpu__time_duration.sum":
            v = float(d["Metric Value"].replace(",", ""))
            u = d["Metric Unit"]
            v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}.get(u, 1.0)
            items.append((re.sub(r"\(.*", "", d["Kernel Name"])[:72], v))
if last:
    items = items[-last:]
agg, cnt = collections.Counter(), collections.Counter()
for n, v in items:
    agg[n] += v
    cnt[n] += 1
tot = sum(agg.values())
for n, v in agg.most_common(30):
    print(f"{v / 1e3:10.1f} us {cnt[n]:5d} {100 * v / tot:5.1f}%  {n}")
print(f"total {tot / 1e3:.1f} us over {len(items)} launches")
