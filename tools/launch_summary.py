"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/launch_summary.py launches.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[ci["Metric Name"]] != "gpu__time_duration.sum":
        continue
    k = r[ci["Kernel Name"]].replace("insb200::", "")
    k = k.split("(CUtensorMap")[0].split("(insb200")[0][:100]
    t = float(r[ci["Metric Value"]].replace(",", ""))
    u = r[ci["Metric Unit"]]
    t = t / 1000 if u in ("ns", "nsecond") else (t * 1000 if u in ("ms", "msecond") else t)
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += t
tot = sum(v[1] for v in agg.values())
print(f"# {sum(v[0] for v in agg.values())} launches, {tot / 1000:.2f} ms in total (times under ncu: cold caches, serialised)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[0]:4d} x {v[1] / v[0]:9.1f} us  {v[1] / tot * 100:5.1f}%  {k}")
