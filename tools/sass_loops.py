"""Static instruction mix of the loops of every kernel in a cuobjdump -sass listing:
    cuobjdump -sass build/fast.o | python tools/sass_loops.py [name filter]"""
import collections
import re
import sys

txt = sys.stdin.read()
flt = sys.argv[1] if len(sys.argv) > 1 else ""
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n")[0]
    if flt not in name:
        continue
    ins = []
    for line in f.split("\n"):
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m:
            ins.append((int(m.group(1), 16), m.group(2)))
    loops = []
    for addr, t in ins:
        m2 = re.search(r"0x([0-9a-f]+)", t) if "BRA" in t else None
        if m2 and int(m2.group(1), 16) < addr:
            loops.append((int(m2.group(1), 16), addr))
    print(name[:100], "total", len(ins))
    for a, b in sorted(loops, key=lambda x: x[0] - x[1])[:3]:
        ops = collections.Counter()
        for addr, t in ins:
            if a <= addr <= b:
                tt = t.split()
                ops[(tt[1] if tt[0].startswith("@") else tt[0]).split(".")[0]] += 1
        print("   loop", hex(a), hex(b), sum(ops.values()), dict(ops.most_common(16)))
