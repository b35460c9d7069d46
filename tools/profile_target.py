#!/usr/bin/env python
"""Small fixed workload for ncu: a few sweeps of T0-v1 (difficulty 5.0) with one kernel variant.

    python tools/profile_target.py [variant] [log2_nonces] [repeats] [version] [sweep|min]

`min` runs upow_search_min over the same range at the unreachable difficulty 16.0 instead (the MODE_MIN kernel
with its limit polls, one launch per repeat).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import hashlib
import upow_b200
from upow_b200 import template as T

variant = int(sys.argv[1]) if len(sys.argv) > 1 else -1
log2n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
version = int(sys.argv[4]) if len(sys.argv) > 4 else 1
mode = sys.argv[5] if len(sys.argv) > 5 else "sweep"
h = lambda b: hashlib.sha256(b).hexdigest()
prev = h(b"prev-block")
wallet = (h(b"addr-x") + h(b"addr-y")) if version == 1 else "DZbXJHETzx7yVQoAZBzt2zdq26c2aP6nfoS6jsR9ZCd3T"
prefix = T.build_block_prefix(prev, T.string_to_bytes(wallet), T.calculate_merkle_root([h(b"tx0"), h(b"tx1"), h(b"tx2")]),
                              1_700_000_000, 5.0)
with upow_b200.Context(1) as ctx:
    if variant >= 0:
        ctx.variant = variant
    ctx.set_template(0, prefix, bytes.fromhex(prev), 5, 16)
    if mode == "min":
        p16 = T.build_block_prefix(prev, T.string_to_bytes(wallet), T.calculate_merkle_root([h(b"tx0"), h(b"tx1"), h(b"tx2")]),
                                   1_700_000_000, 16.0)
        ctx.set_template(1, p16, bytes.fromhex(prev), 16, 16)
        for _ in range(reps):
            rc, res = ctx.search_min(1, max_nonce=1 << log2n)
            assert rc == 1
            print(f"variant {ctx.variant} MIN: {res.hashes_done} hashes, {res.kernel_ms:.3f} ms, "
                  f"{res.hashes_done / res.kernel_ms / 1e6:.2f} GH/s, sm {res.sm_clock_mhz:.0f} MHz")
        reps = 0
    for _ in range(reps):
        nonces, st = ctx.sweep(0, 0, 1 << log2n)
        print(f"variant {ctx.variant} ({ctx.variants()[ctx.variant]}): {len(nonces)} nonces, {st.hashes_done} hashes, "
              f"{st.kernel_ms:.3f} ms, {st.hashes_done / st.kernel_ms / 1e6:.2f} GH/s, sm {st.sm_clock_mhz:.0f} MHz")
