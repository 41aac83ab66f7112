"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the LAST pass that starts
at a kernel whose name contains --start (default: the generator's first conv = start of a training step)."""
import argparse
import collections
import csv
import re

ap = argparse.ArgumentParser()
ap.add_argument("csv")
ap.add_argument("--start", default="first_conv_umma")
ap.add_argument("--top", type=int, default=60)
ap.add_argument("--last", type=int, default=0, help="take the last N launches instead")
a = ap.parse_args()
with open(a.csv) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = list(csv.DictReader(lines))
names = [x["Kernel Name"] for x in rows]
idx = [i for i, n in enumerate(names) if a.start in n]
step = rows[-a.last:] if a.last else (rows[idx[-1]:] if idx else rows)
agg, tot = collections.OrderedDict(), 0.0
for x in step:
    v = float(x["Metric Value"].replace(",", ""))
    u = x["Metric Unit"]
    v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
    n = re.sub(r"\(.*", "", x["Kernel Name"]).replace("void ", "").replace("cid::", "")
    tot += v
    e = agg.setdefault(n, [0, 0.0])
    e[0] += 1
    e[1] += v
print(f"launches {len(step)}  total {tot:.1f} us")
for n, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:a.top]:
    print(f"{v:10.1f} us {100 * v / tot:5.1f}% {c:4d}  {n[:100]}")
