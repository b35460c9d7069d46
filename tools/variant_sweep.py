#!/usr/bin/env python
"""Times every kernel variant of the library on one GPU and checks each against golden data.

    UPOW_ALL_VARIANTS=1 python -m upow_b200.build --force     # the shipped library carries only the plain and the default build
    python tools/variant_sweep.py [--full] [--out gpurun_out/variants.json]

Per variant: (1) parity -- a 2^28-nonce sweep of T0-v1 at difficulty 5.0 against the golden set,
the cfg-2 minimum (6.3), and a v2 sweep against the C oracle; (2) timing -- full 2^32 sweeps of
both layouts, kernel time from CUDA events, SM clock from the kernel's own clock64/globaltimer.
"""
import argparse
import json
import os
import struct
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "variants.json"))
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    import upow_b200
    from oracle import c_oracle
    from oracle import upow_oracle as O

    raw = open(os.path.join(ROOT, "tests", "golden", "sweep_t0v1_d5.u32"), "rb").read()
    golden = list(struct.unpack(f"<{len(raw) // 4}I", raw))
    results = []
    with upow_b200.Context(1) as ctx:
        names = ctx.variants()
        prev = bytes.fromhex(O.T0_PREV_HASH)
        p1, p2 = O.t0_prefix(5.0, 1), O.t0_prefix(5.0, 2)
        want_v2, _ = c_oracle.all_valid_nonces(p2, O.Rule(5.0, O.T0_PREV_HASH), 0, 1 << 24)
        only = [int(x) for x in args.only.split(",") if x]
        for v, name in enumerate(names):
            if only and v not in only:
                continue
            ctx.variant = v
            ctx.set_template(0, p1, prev, 5, 16)
            ctx.set_template(1, p2, prev, 5, 16)
            ctx.set_template(2, O.t0_prefix(6.3), prev, 6, 12)
            got, _ = ctx.sweep(0, 0, 1 << 28)
            ok = got == [n for n in golden if n < (1 << 28)]
            rc, res = ctx.search_min(2)
            ok = ok and (rc, res.nonce) == (0, 9436117)
            got2, _ = ctx.sweep(1, 0, 1 << 24)
            ok = ok and got2 == want_v2
            t1, t2, mhz = [], [], []
            for _ in range(args.reps):
                g, st = ctx.sweep(0, 0, 1 << 32)
                ok = ok and g == golden
                t1.append(st.kernel_ms); mhz.append(st.sm_clock_mhz)
                g, st = ctx.sweep(1, 0, 1 << 32)
                t2.append(st.kernel_ms)
            r = {"variant": v, "name": name, "parity": ok, "grid": ctx.device_info(0).search_grid_blocks,
                 "v1_ms": min(t1), "v1_ghs": 2 ** 32 / min(t1) / 1e6, "v2_ms": min(t2), "v2_ghs": 2 ** 32 / min(t2) / 1e6,
                 "sm_mhz": max(mhz)}
            results.append(r)
            print(f"[{v}] {name:36s} parity={'OK' if ok else 'FAIL'} grid={r['grid']} v1 {r['v1_ghs']:.2f} GH/s "
                  f"({r['v1_ms']:.1f} ms)  v2 {r['v2_ghs']:.2f} GH/s  sm {r['sm_mhz']:.0f} MHz", flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(results, f, indent=1)
    if not all(r["parity"] for r in results):
        sys.exit(1)


if __name__ == "__main__":
    main()
