#!/usr/bin/env python
"""Stress of the launch pipeline, the watchdog and the sticky cancel: a few thousand short searches, sweeps and
batches on one and on three device slots, with cancels and deadlines arriving at random moments from another
thread.  Every completed call must return the oracle's answer; every interrupted call must say so and leave the
context usable.  Run under `timeout` (a hang here would be a bug)."""
import hashlib
import os
import random
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200
from upow_b200 import template as T

h = lambda b: hashlib.sha256(b).hexdigest()
PREV = h(b"prev-block")


def valid(p, rule, n):
    dg = hashlib.sha256(p + n.to_bytes(4, "little")).hexdigest()
    return dg.startswith(PREV[-rule.n_int:]) and (rule.charset_count == 16 or int(dg[rule.n_int], 16) < rule.charset_count)


def cpu_min(p, rule, hi=1 << 32):
    n = 0
    while n < hi:
        if valid(p, rule, n):
            return n
        n += 1
    return None


def main():
    seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = random.Random(12)
    stats = {"ok": 0, "cancelled": 0, "deadline": 0, "batch": 0, "sweep": 0}
    t_end = time.time() + seconds
    configs = [[0], [0, 0, 0]]
    if os.environ.get("UPOW_STRESS_DEVICES"):      # e.g. "0,1": physical GPUs, finds through NVLink peer memory
        configs = [[int(x) for x in os.environ["UPOW_STRESS_DEVICES"].split(",")]] * 2
    for ids in configs:
        with upow_b200.Context(device_ids=ids) as ctx:
            t_ctx_end = time.time() + seconds / 2
            while time.time() < t_ctx_end:
                kind = rng.random()
                d = rng.choice([2.0, 2.5, 3.0, 3.3])
                wallet = h(rng.randbytes(8)) + h(rng.randbytes(8)) if rng.random() < 0.7 else "DZbXJHETzx7yVQoAZBzt2zdq26c2aP6nfoS6jsR9ZCd3T"
                ts = rng.randrange(1, 1 << 31)
                p = T.build_block_prefix(PREV, T.string_to_bytes(wallet), h(b"m"), ts, d)
                rule = T.DifficultyRule.derive(d, PREV)
                if kind < 0.55:      # short search, undisturbed
                    ctx.set_template(0, p, rule.prev_hash, rule.n_int, rule.charset_count)
                    rc, res = ctx.search_min(0)
                    assert rc == 0 and res.nonce == cpu_min(p, rule), (d, ts)
                    stats["ok"] += 1
                elif kind < 0.8:     # long search, cancelled or timed out at a random moment (also before it starts)
                    p16 = T.build_block_prefix(PREV, T.string_to_bytes(wallet), h(b"m"), ts, 16.0)
                    ctx.set_template(1, p16, rule.prev_hash, 16, 16)
                    if rng.random() < 0.5:
                        delay = rng.choice([0.0, 0.0002, 0.001, 0.01, 0.05])
                        timer = threading.Timer(delay, ctx.cancel) if delay else None
                        if timer:
                            timer.start()
                        else:
                            ctx.cancel()
                        rc, res = ctx.search_min(1, max_rollovers=1000)
                        assert rc == upow_b200.UPOW_CANCELLED, rc
                        if timer:
                            timer.join()
                        ctx.clear_cancel()
                        stats["cancelled"] += 1
                    else:
                        rc, res = ctx.search_min(1, max_rollovers=1000, time_limit_s=rng.choice([0.0005, 0.005, 0.03]))
                        assert rc == upow_b200.UPOW_EXHAUSTED and res.found == 0
                        stats["deadline"] += 1
                elif kind < 0.9:     # small batch
                    temps = []
                    for i in range(rng.randrange(1, 20)):
                        pi = T.build_block_prefix(PREV, T.string_to_bytes(wallet), h(b"m%d" % i), ts, d)
                        temps.append((pi, rule.prev_hash, rule.n_int, rule.charset_count))
                    items, _ = ctx.search_batch(temps)
                    k = rng.randrange(len(temps))
                    assert items[k][0] and items[k][1] == cpu_min(temps[k][0], rule)
                    stats["batch"] += 1
                else:                # sweep of a random window
                    ctx.set_template(2, p, rule.prev_hash, rule.n_int, rule.charset_count)
                    lo = rng.randrange(0, (1 << 32) - (1 << 15))
                    got, _ = ctx.sweep(2, lo, lo + (1 << 14))
                    assert got == [n for n in range(lo, lo + (1 << 14)) if valid(p, rule, n)]
                    stats["sweep"] += 1
    print("stress ok:", stats)


if __name__ == "__main__":
    main()
