import ctypes
import json
import os
import struct
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def has_gpu() -> bool:
    try:
        cuda = ctypes.CDLL("libcuda.so.1")
    except OSError:
        return False
    n = ctypes.c_int(0)
    return cuda.cuInit(0) == 0 and cuda.cuDeviceGetCount(ctypes.byref(n)) == 0 and n.value > 0


@pytest.fixture(scope="session")
def mine_block_vectors():
    with open(os.path.join(GOLDEN, "mine_block_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def verifier_vectors():
    with open(os.path.join(GOLDEN, "verifier_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def sweep_fixture():
    with open(os.path.join(GOLDEN, "sweep_t0v1_d5.json")) as f:
        meta = json.load(f)
    raw = open(os.path.join(GOLDEN, "sweep_t0v1_d5.u32"), "rb").read()
    meta["nonces"] = list(struct.unpack(f"<{len(raw) // 4}I", raw))
    return meta


@pytest.fixture(scope="session")
def host_harness():
    """g++ build of tests/host_harness.cpp (the kernels' per-nonce arithmetic on the CPU)."""
    out_dir = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "host_harness.so")
    src = os.path.join(ROOT, "tests", "host_harness.cpp")
    deps = [src] + [os.path.join(ROOT, "upow_b200", "csrc", f) for f in ("template_prep.h", "sha256_core.h", "hot_loop.h")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, src])
    return ctypes.CDLL(so)


@pytest.fixture(scope="session")
def gpu_ctx():
    import upow_b200
    ctx = upow_b200.Context(1)
    yield ctx
    ctx.close()
