#!/usr/bin/env python
"""Condenses an `ncu --set full` report (raw + source CSV pages) into the markdown kept under profiles/.

    ncu -i rep.ncu-rep --page raw --csv > raw.csv; ncu -i rep.ncu-rep --page source --csv > src.csv
    python tools/ncu_summary.py raw.csv src.csv "title" > profiles/xxx.md
"""
import collections
import csv
import sys

raw, src, title = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(open(raw)))
m = {h: (u, v) for h, u, v in zip(rows[0], rows[1], rows[2])}
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "sm__warps_active.avg.per_cycle_active",
        "smsp__issue_active.avg.per_cycle_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum"]
print(f"# {title}\n\n| metric | value |\n|---|---|")
for k in keys:
    if k in m:
        print(f"| `{k}` | {m[k][1]} {m[k][0]} |")
rows = list(csv.reader(open(src)))
print(f"\nKernel: `{rows[0][1]}`\n\n## stall samples per opcode (source page)\n")
print("| opcode | executed (warp inst) | samples | dispatch / issue | math-throttle / issue | wait / issue |\n|---|---|---|---|---|---|")
hdr, data = rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def g(r, k):
    try:
        return float(r[ix[k]] or 0)
    except Exception:
        return 0.0


by = collections.defaultdict(collections.Counter)
for r in data:
    s = r[ix["Source"]].split()
    if not s:
        continue
    op = s[1] if s[0].startswith("@") else s[0]
    op = "IMAD.WIDE" if "IMAD.WIDE" in op else op.split(".")[0]
    for k in ("# Samples", "stall_dispatch", "stall_math", "stall_wait", "stall_selected", "Instructions Executed"):
        by[op][k] += g(r, k)
for op, c in sorted(by.items(), key=lambda x: -x[1]["# Samples"])[:7]:
    sel = c["stall_selected"] or 1
    print(f"| {op} | {c['Instructions Executed']:.3g} | {c['# Samples']:.0f} | {c['stall_dispatch'] / sel:.2f} | "
          f"{c['stall_math'] / sel:.2f} | {c['stall_wait'] / sel:.2f} |")
