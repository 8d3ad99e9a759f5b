"""Prints selected metrics from an ncu report: python scripts/ncu_summary.py report.ncu-rep [regex ...]"""
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
pats = sys.argv[2:] or [
    r"gpu__time_duration\.sum", r"dram__bytes_(read|write)\.sum$", r"gpu__dram_throughput.avg.pct", r"launch__registers_per_thread",
    r"sm__throughput.avg.pct", r"sm__warps_active.avg.pct", r"smsp__issue_active.avg.pct", r"sm__inst_executed_pipe_fp64",
    r"sm__pipe_fp64_cycles_active", r"sm__inst_executed_pipe_(lsu|alu|fma|fmaheavy|uniform|xu|tensor)\b.*sum$", r"smsp__inst_executed.sum$",
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum", r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$",
    r"smsp__average_warps?_issue_stalled.*_per_issue_active", r"smsp__average_warp_latency", r"launch__occupancy_limit",
    r"sm__cycles_elapsed.max", r"smsp__cycles_active.avg$", r"lts__t_bytes.sum$", r"sm__warps_active",
]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("== kernel:", r[hdr.index("Kernel Name")][:60], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for h, u, v in zip(hdr, units, r):
        if any(re.search(p, h) for p in pats):
            print("  %-95s %-14s %s" % (h, u, v))
