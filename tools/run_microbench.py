#!/usr/bin/env python
"""Runs every opcode microbenchmark of the library on GPU 0 and prints / saves the rates."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200

out = upow_b200.microbench(0, iters=1500)
for r in out:
    print(f"{r['name']:36s} {r['warp_inst_per_clk_per_sm']:.3f} warp-inst/clk/SM  (IPC/SMSP {r['warp_inst_per_clk_per_sm'] / 4:.3f})"
          f"  fma share {r['fma_fraction']:.2f}  sm {r['sm_clock_mhz']:.0f} MHz")
path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "microbench.json")
os.makedirs(os.path.dirname(path), exist_ok=True)
json.dump(out, open(path, "w"), indent=1)
