#!/usr/bin/env python
"""GH/s of the batch engine on 4096 templates of one layout at an unreachable difficulty in a bounded nonce
space (fixed work: every template is searched to max_nonce), next to the single-template sweep of the layout."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
import upow_b200
from upow_b200 import template as T

out = {}
with upow_b200.Context(1) as ctx:
    for version in (1, 2):
        temps = []
        for i in range(4096):
            if version == 1:
                p, rule = bench.t0_prefix_and_rule(16.0, bench.T0_TIMESTAMP + i)
            else:
                prev = bench._h(b"prev-block")
                p = T.build_block_prefix(prev, T.string_to_bytes("DZbXJHETzx7yVQoAZBzt2zdq26c2aP6nfoS6jsR9ZCd3T"),
                                         T.calculate_merkle_root([bench._h(b"tx0")]), bench.T0_TIMESTAMP + i, 16.0)
                rule = T.DifficultyRule.derive(16.0, prev)
            temps.append((p, rule.prev_hash, rule.n_int, rule.charset_count))
        arr = ctx.template_array(temps)
        ctx.search_batch(arr, max_nonce=1 << 20)
        best = 0
        for _ in range(2):
            items, st = ctx.search_batch(arr, max_nonce=1 << 24)      # 4096 x 2^24 = 6.9e10 hashes
            assert st.hashes_done == 4096 << 24 and not any(f for f, _, _ in items)
            best = max(best, st.hashes_done / (st.kernel_ms / 1e3) / 1e9)
        ctx.set_template(0, temps[0][0], temps[0][1], 5, 16)
        _, sw = ctx.sweep(0, 0, 1 << 32)
        out[f"v{version}"] = {"batch_ghs": best, "sweep_ghs": sw.hashes_done / (sw.kernel_ms / 1e3) / 1e9}
        print(version, out[f"v{version}"], flush=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "batch_rate.json"), "w"), indent=1)
