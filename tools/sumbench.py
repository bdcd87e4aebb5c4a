"""Print the per-kernel table of bench.py JSON lines: python tools/sumbench.py file.json [...]"""
import json
import sys

for f in sys.argv[1:]:
    try:
        lines = [l for l in open(f).read().splitlines() if l.startswith("{")]
        d = json.loads(lines[-1])
    except Exception as e:  # noqa: BLE001
        print(f, "ERR", e)
        continue
    print(f"{f}: {d['ms_per_step']:.3f} ms/step  value {d['value']:.0f}  e2e {d['e2e']['value']:.0f}  launches {d.get('gpu_launches')}")
    for k, v in sorted(d["kernels"].items(), key=lambda kv: -kv[1]["ms_per_step"]):
        fr = v.get("frac")
        print(f"    {k:24s} {v['ms_per_step']:7.3f} ms/step  {v['launches_per_step']:5.1f} x {v['ms_per_launch']:6.3f} ms  frac {fr if fr is None else round(fr, 3)}")
