#!/usr/bin/env python
"""Writes profiles/sass_counts.json: SASS instructions per hash (inner per-nonce loop) of every
search-kernel variant in the built library, split by issue pipe.  bench.py reads it for the roofline."""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sass_count

lib = os.path.join(ROOT, "upow_b200", "libupow_b200.so")
res = sass_count.analyse(lib, r"search_kernel")
names = subprocess.run(["cu++filt"], input="\n".join(res), capture_output=True, text=True).stdout.split("\n")
kernels = {}
for (mangled, r), dem in zip(res.items(), names):
    p = r["pipes"]
    kernels[dem.strip()] = {"total_per_hash": r["total"], "alu_per_hash": p.get("alu", 0), "fma_per_hash": p.get("fma", 0),
                            "uniform_per_hash": p.get("uniform", 0), "other_per_hash": p.get("other", 0),
                            "loop": r["loop"], "ops": r["ops"]}
default = sys.argv[1] if len(sys.argv) > 1 else None
out = {"generated_by": "tools/make_sass_counts.py (cuobjdump -sass of upow_b200/libupow_b200.so; per-nonce inner loop, 1 nonce per trip)",
       "bench_kernel": default, "kernels": kernels, "dram_bytes_per_launch": None}
if default is None:
    # the default variant's v1 ENUM kernel = what bench.py times
    # default variant = Tune<ADD=1, SHR=0, RR=0, RQ=0, SR=0, SQ=0, CH=0, SHRQ=4>, 3 blocks/SM; layout 1, MODE_ENUM
    cands = [k for k in kernels if "Tune<(int)1, (int)0, (int)0, (int)0, (int)0, (int)0, (int)0, (int)4>, (int)3, (int)256" in k
             and "search_kernel_tuned<(int)1, (int)1," in k]
    out["bench_kernel"] = cands[0] if cands else None
path = os.path.join(ROOT, "profiles", "sass_counts.json")
old = {}
if os.path.exists(path):
    try:
        old = json.load(open(path))
    except Exception:
        pass
out["dram_bytes_per_launch"] = old.get("dram_bytes_per_launch")
json.dump(out, open(path, "w"), indent=1)
for k, v in kernels.items():
    print(f"{v['total_per_hash']:5d} total  {v['alu_per_hash']:5d} alu  {v['fma_per_hash']:4d} fma  {k[:150]}")
print("bench kernel:", out["bench_kernel"])
