"""Summarises an ncu launch list (--metrics gpu__time_duration.sum --csv) into per-kernel shares.
    python tools/launch_shares.py gpurun_out/launches.csv "title" > profiles/rN/launch_shares_x.csv"""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if l.startswith('"')]
rows = list(csv.DictReader(lines))
agg = collections.OrderedDict()
for r in rows:
    n = r["Kernel Name"].split("(")[0]
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += float(r["Metric Value"].replace(",", "")) / 1e3
tot = sum(a[1] for a in agg.values())
print("# " + (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]))
print("kernel,launches,total_us,avg_us,share")
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%s,%d,%.1f,%.2f,%.3f" % (n, a[0], a[1], a[1] / a[0], a[1] / tot))
