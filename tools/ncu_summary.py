"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel-family totals of the LAST step.
usage: python tools/ncu_summary.py launches.csv launches_per_step [out.md]"""
import collections
import csv
import re
import sys

path, per_step = sys.argv[1], int(sys.argv[2])
rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0] != "ID"]
rows = rows[-per_step:]
agg = collections.OrderedDict()
for r in rows:
    name = re.sub(r"^void ", "", r[4])
    name = re.sub(r"\(.*", "", name)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[-1]) / 1e3
tot = sum(a[1] for a in agg.values())
lines = [f"launches {len(rows)}  sum of kernel durations {tot:.1f} us", "", "| launches | total us | share | kernel |", "|---|---|---|---|"]
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    lines.append(f"| {a[0]} | {a[1]:.1f} | {100 * a[1] / tot:.1f}% | `{k}` |")
out = "\n".join(lines)
print(out)
if len(sys.argv) > 3:
    open(sys.argv[3], "w").write(out + "\n")
