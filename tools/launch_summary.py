"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals of the
LAST of `--batches` equal repetitions in the file.  usage: launch_summary.py file.csv [batches]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
data = [(r[ki], float(r[vi].replace(",", ""))) for r in rows[1:]]
batches = int(sys.argv[2]) if len(sys.argv) > 2 else 2
last = data[-(len(data) // batches):]
agg, cnt = collections.defaultdict(float), collections.Counter()
for k, v in last:
    k = k.replace("void ", "").replace("<unnamed>::", "").replace("qre::", "").split("(")[0]
    agg[k] += v
    cnt[k] += 1
tot = sum(agg.values())
for k, v in sorted(agg.items(), key=lambda x: -x[1]):
    print(f"{k:44s} {cnt[k]:4d} x {v / cnt[k] / 1e3:8.1f} us = {v / 1e3:9.1f} us {v / tot:6.1%}")
print(f"{'total':44s} {tot / 1e3:27.1f} us")
