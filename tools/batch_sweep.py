#!/usr/bin/env python
"""cfg 5 (4096 templates at 6.0) with different item lengths / round sizes of the batch engine:
time to solution, GH/s over all evaluated hashes, and the share of hashes beyond the answers."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys, time, json
sys.path.insert(0, %r)
import bench, upow_b200
t = bench.cfg5_templates()
with upow_b200.Context(1) as ctx:
    ctx.search_batch(t[:64])
    best = None
    for _ in range(2):
        r = bench.run_cfg5(ctx, 1, t)
        if best is None or r["wall_s"] < best["wall_s"]:
            best = r
print(json.dumps(best))
""" % ROOT
out = []
for hc in (16, 32, 64):
    for poll in (1, 2, 4, 8):
        rnd = 0
        env = dict(os.environ, UPOW_BATCH_HC=str(hc), UPOW_BATCH_POLL=str(poll))
        r = json.loads(subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1])
        out.append({"hc": hc, "poll": poll, "wall_s": r["wall_s"], "ghs": r["ghs"], "overshoot": r["overshoot"]})
        print(out[-1], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "batch_items.json"), "w"), indent=1)
