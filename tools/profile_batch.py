#!/usr/bin/env python
"""Small fixed batch workload for ncu: 256 templates (SURVEY 8(d) cfg-5 derivation) at difficulty 6.0."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200
from upow_b200 import template as T

h = lambda b: hashlib.sha256(b).hexdigest()
PREV = h(b"prev-block")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
rule = T.DifficultyRule.derive(6.0, PREV)
temps = []
for i in range(n):
    le = i.to_bytes(4, "little")
    addr = h(b"addr-x" + le) + h(b"addr-y" + le)
    txs = [h(b"tx" + le + bytes([j])) for j in range(3)]
    temps.append((T.build_block_prefix(PREV, T.string_to_bytes(addr), T.calculate_merkle_root(txs), 1_700_000_000, 6.0),
                  rule.prev_hash, rule.n_int, rule.charset_count))
with upow_b200.Context(1) as ctx:
    items, st = ctx.search_batch(temps)
    print(f"{n} templates: {sum(f for f, _, _ in items)} found, {st.hashes_done} hashes, {st.kernel_ms:.2f} ms, "
          f"{st.hashes_done / st.kernel_ms / 1e6:.2f} GH/s, {st.launches} launches")
