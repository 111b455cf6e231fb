"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python profiles/launch_summary.py gpurun_out/launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r and r[0].isdigit()]
d = collections.defaultdict(list)
for r in rows:
    d[r[4].split("(")[0].replace("void ", "").replace("b2::", "").replace("<unnamed>::", "").replace("unnamed>::", "")].append(float(r[-1]))
tot = sum(sum(v) for v in d.values())
print(f"{'kernel':44s} {'n':>5s} {'mean us':>10s} {'sum us':>12s} {'share':>7s}")
for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
    print(f"{k:44s} {len(v):5d} {sum(v) / len(v) / 1e3:10.2f} {sum(v) / 1e3:12.1f} {100 * sum(v) / tot:6.1f}%")
