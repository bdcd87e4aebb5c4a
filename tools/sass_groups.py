"""Group the SASS of one kernel in an `ncu --page source --csv` export by execution count.
usage: python tools/sass_groups.py src.csv START_LINE [END_LINE] [--stalls]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
a = int(sys.argv[2]) - 1
b = int(sys.argv[3]) - 1 if len(sys.argv) > 3 and sys.argv[3].isdigit() else len(rows)
rows = rows[a:b]
hi = [i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r][0]
hdr = rows[hi]
ci = {h: i for i, h in enumerate(hdr)}
src, ex, st = ci["Source"], ci["Instructions Executed"], ci["Warp Stall Sampling (All Samples)"]
groups, tot, stot = [], 0, 0
for r in rows[hi + 1:]:
    try:
        n = float(r[ex])
    except (ValueError, IndexError):
        continue
    s = r[src].strip()
    toks = s.split()
    op = (toks[1] if toks[0].startswith("@") else toks[0]).split(".")[0]
    tot += n
    sv = float(r[st] or 0)
    stot += sv
    if groups and abs(groups[-1][0] - n) <= 0.02 * max(n, 1):
        groups[-1][1][op] += 1
        groups[-1][2] += n
        groups[-1][3] += sv
    else:
        groups.append([n, collections.Counter({op: 1}), n, sv])
print("total warp instructions", tot)
for n, c, sm, sv in groups:
    t = sum(c.values())
    if sm > 0.004 * tot:
        print(f"{n:12.0f} x {t:4d} = {sm:.3g} ({sm / tot * 100:4.1f}% instr, {sv / max(stot, 1) * 100:4.1f}% stall samples) ", dict(c.most_common(8)))
