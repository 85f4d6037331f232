#!/usr/bin/env python3
"""Digest of one kernel in an .ncu-rep: headline metrics, stall reasons per issue, and the SASS lines that
collect the most stall samples.   python tools/ncu_digest.py report.ncu-rep [n_top]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 8


def page(*args):
    out = subprocess.run(["ncu", "-i", rep, "--csv", *args], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


raw = page("--page", "raw")
hdr, units, vals = raw[0], raw[1], raw[2]
want = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem"]
for i, h in enumerate(hdr):
    if h in want:
        print(f"{h} = {vals[i]} {units[i]}")
st = [(float(vals[i]), h) for i, h in enumerate(hdr)
      if "average_warps_issue_stalled" in h and h.endswith("_per_issue_active.ratio")]
print("stalls per issue:", ", ".join(f"{h.split('stalled_')[1].split('_per_')[0]} {v:.2f}" for v, h in sorted(st, reverse=True)[:9]))
src = page("--page", "source", "--print-source", "sass")
h2 = src[1]
ix = {h: i for i, h in enumerate(h2)}
data = src[2:]
base = min(int(r[ix["Address"]], 16) for r in data)
for col in ("stall_long_sb", "stall_barrier", "stall_short_sb", "stall_math", "stall_mio", "stall_wait", "stall_not_selected"):
    tot = sum(int(r[ix[col]] or 0) for r in data)
    top = sorted(data, key=lambda r: -int(r[ix[col]] or 0))[:ntop]
    print(f"== {col} total {tot}")
    for r in top:
        if int(r[ix[col]] or 0) * 50 > tot:
            print(f"   +{int(r[ix['Address']], 16) - base:05x} {r[ix[col]]:>6}  {r[ix['Source']].strip()[:80]}")
