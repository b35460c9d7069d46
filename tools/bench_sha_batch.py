#!/usr/bin/env python
"""Throughput of upow_sha256_batch: messages/s, input GB/s vs the measured HBM copy peak,
SHA-256 compressions/s vs what the search kernel reaches, and hashlib on one host core."""
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

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
size = int(sys.argv[2]) if len(sys.argv) > 2 else 250          # a typical transaction blob
data = os.urandom(1 << 20) * (n * size // (1 << 20) + 1)
data = data[: n * size]
offs = (C.c_uint64 * (n + 1))(*range(0, n * size + 1, size))
out = C.create_string_buffer(32 * n)
with upow_b200.Context(1) as ctx:
    stats = _lib.Result()
    best, e2e = None, None
    for _ in range(4):
        rc = ctx._L.upow_sha256_batch(ctx._h, n, data, offs, out, C.byref(stats))
        assert rc == 0
        best = min(best or 1e9, stats.kernel_ms)
        e2e = min(e2e or 1e9, stats.elapsed_s)
for i in range(0, n, n // 50):
    assert out.raw[32 * i: 32 * i + 32] == hashlib.sha256(data[i * size:(i + 1) * size]).digest()
t = time.perf_counter()
m = 200_000
for i in range(m):
    hashlib.sha256(data[i * size:(i + 1) * size]).digest()
cpu = m / (time.perf_counter() - t)
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except OSError:
    pass
blocks = (size + 9 + 63) // 64
gbs = n * (size + 32) / (best / 1e3) / 1e9
res = {"messages": n, "bytes_each": size, "kernel_ms": best, "messages_per_s_kernel": n / (best / 1e3),
       "compressions_per_s": n * blocks / (best / 1e3), "hbm_gbs_algorithmic": gbs, "hbm_peak_gbs": peaks.get("hbm_gbs"),
       "frac_of_measured_hbm_peak": gbs / peaks["hbm_gbs"] if peaks.get("hbm_gbs") else None,
       "e2e_messages_per_s": n / e2e, "cpu_messages_per_s_one_core": cpu}
print(json.dumps(res))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "sha_batch_bench.json"), "w"), indent=1)
