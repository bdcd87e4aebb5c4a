"""Summarise an .ncu-rep: per launch duration, DRAM bytes, executed warp instructions, issue utilisation, shared-memory
wavefronts; optionally the executed instruction mix per warp-plane of one launch.
    python tools/ncu_summary.py file.ncu-rep [cells]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
cells = float(sys.argv[2]) if len(sys.argv) > 2 else 134217728.0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
wp = cells / 32
for r in rows[2:]:
    name = r[idx["Kernel Name"]]
    print(name[:90])
    for c in cols:
        if c in idx:
            v = r[idx[c]]
            extra = ""
            if c in ("smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"):
                extra = f"   = {float(v) / wp:.1f} per warp-plane (32 cells)"
            print(f"    {c:72s} {v:>16s} {units[idx[c]]}{extra}")
