#!/usr/bin/env python
"""BASELINE configs 4 and 5 at full size on the GPUs of this box; writes gpurun_out/configs.json.

    python tools/run_configs.py [--gpus N] [--difficulty9 9.0] [--templates 4096]

config 4: T0-v1 at difficulty 9.x, rollover ts <- ts+1 on nonce exhaustion, lexicographic (ts, nonce)
          minimum; run on 1..N GPUs -- every run must return the same (ts, nonce, digest).
config 5: 4096 templates at difficulty 6.0, per-template minimum nonce, all GPUs.
Results are checked for validity with hashlib here; minimality checks live in tests/.
"""
import argparse
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200
from upow_b200 import template as T

h = lambda b: hashlib.sha256(b).hexdigest()
PREV = h(b"prev-block")
T0 = 1_700_000_000


def t0_prefix(difficulty):
    wallet = h(b"addr-x") + h(b"addr-y")
    return T.build_block_prefix(PREV, T.string_to_bytes(wallet), T.calculate_merkle_root([h(b"tx0"), h(b"tx1"), h(b"tx2")]),
                                T0, difficulty)


def batch_prefix(i, difficulty):
    le = i.to_bytes(4, "little")
    addr = h(b"addr-x" + le) + h(b"addr-y" + le)
    txs = [h(b"tx" + le + bytes([j])) for j in range(3)]
    return T.build_block_prefix(PREV, T.string_to_bytes(addr), T.calculate_merkle_root(txs), T0, difficulty)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--difficulty9", type=float, nargs="+", default=[9.0, 9.3, 9.5])
    ap.add_argument("--templates", type=int, default=4096)
    ap.add_argument("--skip4", action="store_true", help="config 5 only")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.json"))
    a = ap.parse_args()
    out = {"gpus": a.gpus, "config4": [], "config5": None}
    for d9 in ([] if a.skip4 else a.difficulty9):
        rule = T.DifficultyRule.derive(d9, PREV)
        prefix = t0_prefix(d9)
        n = 1
        ref = None
        while n <= a.gpus:
            with upow_b200.Context(n) as ctx:
                ctx.set_template(0, prefix, rule.prev_hash, rule.n_int, rule.charset_count)
                t = time.perf_counter()
                rc, res = ctx.search_min(0, max_rollovers=4096)
                wall = time.perf_counter() - t
                block = T.with_timestamp(prefix, res.timestamp) + res.nonce.to_bytes(4, "little")
                digest = hashlib.sha256(block).hexdigest()
                ok = rc == 0 and digest == res.digest_hex and rule.accepts(digest)
                rec = {"n_gpus": n, "difficulty": d9, "rc": rc, "timestamp": res.timestamp, "rollovers": res.rollovers,
                       "nonce": res.nonce, "digest": res.digest_hex, "valid": ok, "hashes": res.hashes_done,
                       "wall_s": wall, "ghs": res.hashes_done / wall / 1e9, "launches": res.launches}
                ref = ref or (res.timestamp, res.nonce, res.digest_hex)
                rec["same_as_1gpu"] = (res.timestamp, res.nonce, res.digest_hex) == ref
                out["config4"].append(rec)
                print(rec, flush=True)
            n *= 2
    d5 = 6.0
    rule5 = T.DifficultyRule.derive(d5, PREV)
    temps = [(batch_prefix(i, d5), rule5.prev_hash, rule5.n_int, rule5.charset_count) for i in range(a.templates)]
    with upow_b200.Context(a.gpus) as ctx:
        t = time.perf_counter()
        items, stats = ctx.search_batch(temps)
        wall = time.perf_counter() - t
    valid = 0
    for (found, nonce, digest), (p, _, _, _) in zip(items, temps):
        dg = hashlib.sha256(p + nonce.to_bytes(4, "little")).hexdigest()
        valid += bool(found and dg == digest.hex() and rule5.accepts(dg))
    fp = hashlib.sha256(b"".join(n.to_bytes(4, "little") for _, n, _ in items)).hexdigest()
    out["config5"] = {"n_gpus": a.gpus, "templates": a.templates, "difficulty": d5, "all_valid": valid == a.templates,
                      "hashes": stats.hashes_done, "wall_s": wall, "ghs": stats.hashes_done / wall / 1e9,
                      "kernel_ms_max": stats.kernel_ms, "launches": stats.launches, "nonce_fingerprint": fp}
    print(out["config5"], flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(out, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
