#!/usr/bin/env python
"""Small pass through every kernel of libupow_b200.so, sized for compute-sanitizer:

    compute-sanitizer --tool memcheck  python tools/sanitizer_target.py
    compute-sanitizer --tool racecheck python tools/sanitizer_target.py
    compute-sanitizer --tool synccheck python tools/sanitizer_target.py

Every result is still checked against hashlib, so a sanitizer run is also a (tiny) parity run.  The find /
limit / kill-word protocol, the hit buffer of the sweep, the one-word overread of the batched hashing kernel
and the shared-memory staging of the block verifier are all exercised, with one and with two device queues."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200
from upow_b200 import _lib
from upow_b200 import template as T

h = lambda b: hashlib.sha256(b).hexdigest()
PREV = h(b"prev-block")


def prefix(d, v=1, ts=1_700_000_000, i=None):
    wallet = (h(b"addr-x") + h(b"addr-y")) if v == 1 else "DZbXJHETzx7yVQoAZBzt2zdq26c2aP6nfoS6jsR9ZCd3T"
    if i is not None:
        le = i.to_bytes(4, "little")
        wallet = h(b"addr-x" + le) + h(b"addr-y" + le)
    txs = [h(b"tx0"), h(b"tx1"), h(b"tx2")]
    return T.build_block_prefix(PREV, T.string_to_bytes(wallet), T.calculate_merkle_root(txs), ts, d), T.DifficultyRule.derive(d, PREV)


def valid(p, rule, n):
    dg = hashlib.sha256(p + n.to_bytes(4, "little")).hexdigest()
    return dg.startswith(PREV[-rule.n_int:]) and (rule.charset_count == 16 or int(dg[rule.n_int], 16) < rule.charset_count)


def cpu_min(p, rule, start=0, stride=1, hi=1 << 32):
    n = start
    while n < hi:
        if valid(p, rule, n):
            return n
        n += stride
    return None


def main():
    assert upow_b200.selftest(0) == 0
    for ids in ([0], [0, 0]):
        with upow_b200.Context(device_ids=ids) as ctx:
            for v, d in ((1, 3.0), (2, 3.3), (1, 2.0)):
                p, r = prefix(d, v)
                ctx.set_template(0, p, r.prev_hash, r.n_int, r.charset_count)
                rc, res = ctx.search_min(0)
                assert rc == 0 and res.nonce == cpu_min(p, r), (ids, v, d)
                rc, res = ctx.search_min(0, 3, 7)
                assert rc == 0 and res.nonce == cpu_min(p, r, 3, 7)
                rc, res = ctx.search_min(0, max_nonce=300, max_rollovers=40, tile_nonces=1 << 16)
                got, _ = ctx.sweep(0, 1000, 1000 + (1 << 17))
                assert got == [n for n in range(1000, 1000 + (1 << 17)) if valid(p, r, n)]
            ctx.variant = 0
            p, r = prefix(3.0, 1)
            ctx.set_template(1, p, r.prev_hash, r.n_int, r.charset_count)
            rc, res = ctx.search_min(1)
            assert res.nonce == cpu_min(p, r)
            rc, res = ctx.search_min(1, 1, 2)
            assert res.nonce == cpu_min(p, r, 1, 2)
            ctx.variant = ctx.default_variant
            temps = []
            for i in range(12):
                p, r = prefix(2.5, 2 if i % 3 == 0 else 1, 1_700_000_000 + i, None if i % 3 == 0 else i)
                temps.append((p, r.prev_hash, r.n_int, r.charset_count))
            items, _ = ctx.search_batch(temps)
            for (found, nonce, dg), (p, _, n_int, cc) in zip(items, temps):
                assert found and nonce == cpu_min(p, T.DifficultyRule.derive(2.5, PREV)) and dg == hashlib.sha256(p + nonce.to_bytes(4, "little")).digest()
            rc, idx, nonce, dg, _ = ctx.search_any(temps)
            assert rc == 0 and dg == hashlib.sha256(temps[idx][0] + nonce.to_bytes(4, "little")).digest()
            blocks = [(temps[i % 12][0] + (i * 2654435761 % (1 << 32)).to_bytes(4, "little"), temps[i % 12][1], 1, 16) for i in range(300)]
            blocks.append((b"\x01" * 137, bytes(32), 4, 16))
            verdicts, digests, _ = ctx.verify_blocks(blocks)
            for (c, *_), dg, v in zip(blocks, digests, verdicts):
                if v != 255:
                    assert dg == hashlib.sha256(c).digest()
            msgs = [bytes((i * 7 + j) & 255 for j in range(n)) for i, n in enumerate([0, 1, 3, 55, 56, 63, 64, 65, 119, 120, 138, 200, 1000])]
            dgs, _ = ctx.sha256_many(msgs)
            assert dgs == [hashlib.sha256(m).digest() for m in msgs]
    p, r = prefix(3.0, 1)
    got = _lib.sweep_noalloc(0, p, r.prev_hash, r.n_int, r.charset_count, 0, 1 << 17)
    assert got == [n for n in range(1 << 17) if valid(p, r, n)]
    print("sanitizer target: all results match hashlib")


if __name__ == "__main__":
    main()
