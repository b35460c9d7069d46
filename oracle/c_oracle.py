"""ctypes binding of oracle/upow_oracle.c -- TEST INFRASTRUCTURE ONLY (see that file's header)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libupow_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "upow_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE] + (["-B"] if force else []))
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
        L.upow_oracle_sha256.argtypes = [C.c_char_p, C.c_uint64, C.c_char_p]
        L.upow_oracle_tail_digest.argtypes = [C.c_char_p, C.c_int, C.c_uint32, C.c_char_p]
        L.upow_oracle_set_shani.argtypes = [C.c_int]
        L.upow_oracle_min_nonce.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                            C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, u64p, u64p]
        L.upow_oracle_sweep.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                        C.c_uint64, C.c_uint64, C.c_int, u32p, C.c_uint64, u64p, u64p]
        _lib = L
    return _lib


def _rule_args(rule):
    """Nibble form of an upow_oracle.Rule (target string -> nibble bytes)."""
    tgt = bytes(int(ch, 16) for ch in rule.target_prefix)
    return tgt, len(tgt), rule.int_difficulty, rule.charset_count


def sha256(msg: bytes) -> bytes:
    out = C.create_string_buffer(32)
    lib().upow_oracle_sha256(msg, len(msg), out)
    return out.raw


def tail_digest(prefix: bytes, nonce: int) -> bytes:
    out = C.create_string_buffer(32)
    rc = lib().upow_oracle_tail_digest(prefix, len(prefix), nonce, out)
    assert rc == 0
    return out.raw


def set_shani(enable: bool) -> bool:
    return bool(lib().upow_oracle_set_shani(int(enable)))


def min_valid_nonce(prefix: bytes, rule, start_nonce=0, nonce_increment=1, max_nonce=1 << 32, threads=None):
    tgt, n, fi, cc = _rule_args(rule)
    out, hashes = C.c_uint64(0), C.c_uint64(0)
    rc = lib().upow_oracle_min_nonce(prefix, len(prefix), tgt, n, fi, cc, start_nonce, nonce_increment,
                                     max_nonce, threads or os.cpu_count() or 1, C.byref(out), C.byref(hashes))
    if rc < 0:
        raise ValueError("bad arguments")
    return (out.value if rc == 0 else None), hashes.value


def all_valid_nonces(prefix: bytes, rule, lo: int, hi: int, threads=None, cap=1 << 22):
    tgt, n, fi, cc = _rule_args(rule)
    buf = (C.c_uint32 * cap)()
    count, hashes = C.c_uint64(0), C.c_uint64(0)
    rc = lib().upow_oracle_sweep(prefix, len(prefix), tgt, n, fi, cc, lo, hi, threads or os.cpu_count() or 1,
                                 buf, cap, C.byref(count), C.byref(hashes))
    if rc != 0:
        raise RuntimeError(f"oracle sweep rc={rc}")
    return sorted(buf[: count.value]), hashes.value
