"""Per-kernel table of ONE training step from an ncu launch list
(`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv --log-file F ...`).
The step is the span between the last two launches of the marker kernel (default: the optimizer).
    python tools/agg_step.py F [marker-substring] [peak GB/s]"""
import collections
import csv
import re
import sys

path = sys.argv[1]
marker = sys.argv[2] if len(sys.argv) > 2 else "optimizer_kernel"
peak = float(sys.argv[3]) if len(sys.argv) > 3 else 6542.0
UNIT = {"ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6,
        "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

hdr, launches = None, collections.OrderedDict()
for r in csv.reader(open(path, errors="ignore")):
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        try:
            v = float(d["Metric Value"].replace(",", "")) * UNIT.get(d["Metric Unit"], 1.0)
        except ValueError:
            continue
        name = d["Kernel Name"]
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"\((?:[^()]|\([^()]*\))*\)\s*$", "", name).replace("<unnamed>::", "")
        e = launches.setdefault(int(d["ID"]), {"name": name[:80], "t": 0.0, "b": 0.0})
        if d["Metric Name"] == "gpu__time_duration.sum":
            e["t"] = v
        elif d["Metric Name"].startswith("dram__bytes"):
            e["b"] += v
items = list(launches.values())
marks = [i for i, e in enumerate(items) if marker in e["name"]]
if len(marks) >= 2:
    items = items[marks[-2] + 1: marks[-1] + 1]
agg = collections.OrderedDict()
for e in items:
    a = agg.setdefault(e["name"], {"n": 0, "t": 0.0, "b": 0.0, "best": 0.0})
    a["n"] += 1
    a["t"] += e["t"]
    a["b"] += e["b"]
    if e["t"] > 0:
        a["best"] = max(a["best"], e["b"] / e["t"])
tot = sum(a["t"] for a in agg.values())
print("| kernel | launches | total µs | share | avg DRAM GB/s | best-launch DRAM GB/s (fraction of %d) |" % peak)
print("|---|---|---|---|---|---|")
for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
    if a["t"] / tot < 0.003:
        continue
    print(f"| `{name}` | {a['n']} | {a['t'] / 1e3:.0f} | {100 * a['t'] / tot:.1f} % | {a['b'] / a['t']:.0f} | "
          f"{a['best']:.0f} ({a['best'] / peak:.2f}) |")
print(f"\nTotal {tot / 1e6:.2f} ms over {len(items)} launches.")
