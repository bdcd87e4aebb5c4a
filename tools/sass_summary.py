"""Per-kernel SASS evidence of the Blackwell-native path: counts of TMA (UTMALDG / UTMASTG), mbarrier (SYNCS), bulk-copy
and FP instructions in every kernel of the built library.
    python tools/sass_summary.py [lib/libinsb200.so] > profiles/rN_sass_hot_kernels.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "incompressiblenavierstokes.jl_b200", "lib", "libinsb200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
archs = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}: cubin architectures {archs}")
print("# kernel | instructions | UTMALDG | UTMASTG | SYNCS (mbarrier) | UBLKCP | FFMA | FADD | FMUL | DFMA | LDS | STS | LDG | STG | BAR")
keys = ["UTMALDG", "UTMASTG", "SYNCS", "UBLKCP", "FFMA", "FADD", "FMUL", "DFMA", "LDS", "STS", "LDG", "STG", "BAR"]
tot = collections.Counter()
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n")[0].strip()
    dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip() or name
    dem = re.sub(r"\((bool|int)\)", "", dem)
    dem = re.sub(r">\(.*", ">", dem).replace("insb200::", "").replace("void ", "")
    dem = re.sub(r"\(.*", "", dem) if "<" not in dem else dem
    ops = collections.Counter()
    n = 0
    for line in f.split("\n"):
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if not m:
            continue
        t = m.group(1).split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ops[op] += 1
        n += 1
    hot = any(k in dem for k in ("k_stage3d_fast", "k_stage3d_tma<", "k_smag3d_v2", "k_zsolve<", "k_fft_scatter2", "k_peer_barrier", "k_halo_push",
                                 "k_gradient_z", "k_fill_ghosts", "k_stage2d_warp", "k_reduce_partials"))
    for k in keys:
        tot[k] += ops[k]
    if hot:
        print(f"{dem[:92]:92s} | {n:5d} | " + " | ".join(f"{ops[k]:4d}" for k in keys))
print("# whole library: " + ", ".join(f"{k} {tot[k]}" for k in keys))
