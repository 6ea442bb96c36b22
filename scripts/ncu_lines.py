"""Aggregate an ncu source page (cuda,sass) by source line: share of executed instructions and of stall samples.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python scripts/ncu_lines.py src.csv [N]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
inst, stall, src = collections.Counter(), collections.Counter(), {}
for r in rows:
    if len(r) < 8 or not r[0].isdigit():
        continue
    try:
        n, s = int(r[7]), int(r[4])
    except ValueError:
        continue
    key = int(r[0])
    inst[key] += n
    stall[key] += s
    src[key] = r[1].strip()[:100]
ti, ts = sum(inst.values()), max(1, sum(stall.values()))
print("warp instructions:", ti, "stall samples:", ts)
for k, v in inst.most_common(top):
    print("%4d %5.1f%% inst %5.1f%% stall  %s" % (k, 100 * v / ti, 100 * stall[k] / ts, src[k]))
