"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per (kernel, grid)."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "")
    agg[(name, r[8], r[7])][0] += 1
    agg[(name, r[8], r[7])][1] += float(r[-1])
tot = sum(v[1] for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-52s grid %-14s blk %-12s n=%5d avg %8.1f us share %5.1f%%" % (k[0][:52], k[1], k[2], v[0], v[1] / v[0] / 1e3, 100 * v[1] / tot))
print("total %.3f ms over %d launches" % (tot / 1e6, len(rows)))
