#!/usr/bin/env python
"""Throughput of upow_verify_blocks straight through the C-ABI with a preformatted
record array: blocks/s, GB/s of HBM traffic (192 B in + 33 B out per block) vs the measured copy peak, and
the CPU figure for the same work (hashlib + the oracle's rule, one core)."""
import ctypes as C
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import upow_b200
from upow_b200 import _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
with upow_b200.Context(1) as ctx:
    arr = (_lib.Block * n)()
    h = lambda b: hashlib.sha256(b).digest()
    prev = h(b"prev-block")
    content = (prev + h(b"addr-x") + h(b"addr-y") + h(b"m") + (1_700_000_000).to_bytes(4, "little") + (15).to_bytes(2, "little"))
    proto = _lib.Block()
    C.memmove(proto.content, content + b"\0\0\0\0", 138)
    proto.content_len, proto.n_int, proto.charset_count = 138, 1, 8
    C.memmove(proto.prev_hash, prev, 32)
    buf = bytearray(bytes(proto) * n)
    for i in range(0, n):                      # distinct nonce per record
        buf[192 * i + 134: 192 * i + 138] = i.to_bytes(4, "little")
    C.memmove(arr, bytes(buf), len(buf))
    valid = C.create_string_buffer(n)
    stats = _lib.Result()
    best = None
    e2e = 1e9
    for _ in range(4):
        rc = ctx._L.upow_verify_blocks(ctx._h, n, arr, valid, None, C.byref(stats))
        assert rc == 0
        e2e = min(e2e, stats.elapsed_s)
    # kernel alone, records resident on the device (the *_dev entry point)
    import torch
    d_rec = torch.frombuffer(buf, dtype=torch.uint8).cuda()
    d_valid = torch.zeros(n, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(4):
        rc = ctx._L.upow_verify_blocks_dev(ctx._h, 0, n, d_rec.data_ptr(), d_valid.data_ptr(), None, C.byref(stats))
        assert rc == 0
        best = min(best or 1e9, stats.kernel_ms)
    assert bytes(d_valid.cpu().numpy()) == valid.raw
    # what the link gives: pinned host -> device copy of the same bytes
    pin = torch.empty(len(buf), dtype=torch.uint8).pin_memory()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h2d = 0.0
    for _ in range(3):
        ev0.record(); d_rec.copy_(pin, non_blocking=True); ev1.record(); torch.cuda.synchronize()
        h2d = max(h2d, len(buf) / (ev0.elapsed_time(ev1) / 1e3) / 1e9)
    t = time.perf_counter()
    m = 20000
    ok = 0
    for i in range(m):
        d = hashlib.sha256(bytes(buf[192 * i: 192 * i + 138])).hexdigest()
        ok += d.startswith(prev.hex()[-1:]) and d[1] in "01234567"
    cpu = m / (time.perf_counter() - t)
    assert ok == sum(valid.raw[:m])
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    gbs = n * (192 + 1) / (best / 1e3) / 1e9
    out = {"blocks": n, "kernel_ms": best, "blocks_per_s_kernel": n / (best / 1e3), "hbm_gbs_algorithmic": gbs,
           "hbm_peak_gbs": peaks.get("hbm_gbs"), "frac_of_measured_peak": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
           "e2e_blocks_per_s": n / e2e, "e2e_host_gbs": n * 193 / e2e / 1e9, "pinned_h2d_peak_gbs": h2d,
           "e2e_over_round1": (n / e2e) / 5.7e7,
           "cpu_blocks_per_s_one_core": cpu, "valid_share": sum(valid.raw) / n}
    print(json.dumps(out))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "verify_bench.json"), "w"), indent=1)
