#!/usr/bin/env python
"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> markdown share table.
usage: summarize_launches.py launches.csv "title" "command" [live_ms_per_step] > profiles/x.md"""
import collections
import csv
import re
import sys

path, title, cmd = sys.argv[1:4]
live = float(sys.argv[4]) if len(sys.argv) > 4 else None
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", ""))
    v = {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(row["Metric Unit"], v)
    name = re.sub(r"\(.*", "", row["Kernel Name"])[:72]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
n = sum(v[0] for v in agg.values())
print(f"# {title}\n\nCommand: `{cmd}`.\nTimes are cold-cache, serialised: compare shares, not absolutes.\n")
print(f"Total kernel time {tot / 1e3:.1f} ms over {n} launches" +
      (f" (plain run of the same command: {live:.1f} ms/step)." if live else "."))
print("\n| share | total us | launches | avg us | kernel |\n|---|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if v[1] / tot < 0.0005:
        continue
    print(f"| {100 * v[1] / tot:.2f}% | {v[1]:.1f} | {v[0]} | {v[1] / v[0]:.1f} | `{k}` |")
