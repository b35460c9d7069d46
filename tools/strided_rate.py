import sys, os
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import upow_b200
from oracle import upow_oracle as O
with upow_b200.Context(1) as ctx:
    for version in (1, 2):
        r = O.Rule(16.0, O.T0_PREV_HASH)
        ctx.set_template(0, O.t0_prefix(16.0, version), bytes.fromhex(O.T0_PREV_HASH), r.int_difficulty, r.charset_count)
        for v in (0, ctx.variant):
            ctx.variant = v
            rc, res = ctx.search_min(0, 1, 3)
            print(f"v{version} variant {v}: stride 3 full space: {res.hashes_done} hashes, {res.hashes_done/res.kernel_ms/1e6:.2f} GH/s")
        ctx.variant = 8
