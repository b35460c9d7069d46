"""ctypes binding of libupow_b200.so (include/upow_b200.h).  No torch, no CPU fallback:
importing works without a GPU (so the symbol table can be checked), but creating a
``Context`` without a CUDA device raises ``UpowError``.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

UPOW_MAX_DEVICES = 8
UPOW_MAX_SLOTS = 64
UPOW_OK, UPOW_EXHAUSTED, UPOW_CANCELLED = 0, 1, 2
UPOW_E_ARG, UPOW_E_CUDA, UPOW_E_OVERFLOW, UPOW_E_INTERNAL = -1, -2, -3, -4

EXPORTS = [
    "upow_init", "upow_destroy", "upow_device_count", "upow_get_device_info", "upow_last_error",
    "upow_set_template", "upow_search_min", "upow_sweep", "upow_search_batch", "upow_cancel",
    "upow_selftest", "upow_microbench_count", "upow_microbench", "upow_flush_l2", "upow_version",
    "upow_variant_count", "upow_variant_name", "upow_get_variant", "upow_set_variant",
    "upow_sweep_noalloc", "upow_noalloc_last_error", "upow_search_any", "upow_verify_blocks", "upow_sha256_batch",
    "upow_clear_cancel", "upow_get_progress", "upow_default_variant", "upow_verify_blocks_dev", "upow_sha256_batch_dev",
]
NOALLOC_MAX_HITS = 65536


class UpowError(RuntimeError):
    def __init__(self, code: int, where: str, text: str = ""):
        super().__init__(f"{where}: rc={code} {text}".strip())
        self.code = code


class Limits(C.Structure):
    _fields_ = [("max_nonce", C.c_uint64), ("max_rollovers", C.c_uint32), ("rollover_backwards", C.c_uint32),
                ("time_limit_s", C.c_double), ("tile_nonces", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [("found", C.c_int32), ("rollovers", C.c_uint32), ("timestamp", C.c_uint32), ("nonce", C.c_uint32),
                ("digest", C.c_uint8 * 32), ("hashes_done", C.c_uint64), ("elapsed_s", C.c_double),
                ("kernel_ms", C.c_double), ("per_device_hashes", C.c_uint64 * UPOW_MAX_DEVICES),
                ("per_device_kernel_ms", C.c_double * UPOW_MAX_DEVICES), ("sm_clock_mhz", C.c_double),
                ("launches", C.c_uint32), ("reserved0", C.c_uint32)]

    @property
    def digest_hex(self) -> str:
        return bytes(self.digest).hex()


class DeviceInfo(C.Structure):
    _fields_ = [("name", C.c_char * 64), ("device_id", C.c_int32), ("sm_count", C.c_int32), ("cc_major", C.c_int32),
                ("cc_minor", C.c_int32), ("max_sm_clock_khz", C.c_int32), ("search_grid_blocks", C.c_int32),
                ("total_mem", C.c_uint64), ("peer_devices", C.c_int32), ("reserved0", C.c_int32)]


class MicrobenchResult(C.Structure):
    _fields_ = [("name", C.c_char * 48), ("sm_count", C.c_int32), ("reserved0", C.c_int32),
                ("kernel_ms", C.c_double), ("max_block_cycles", C.c_double), ("mean_block_cycles", C.c_double),
                ("warp_inst_per_clk_per_sm", C.c_double), ("lanes_per_clk_per_sm", C.c_double),
                ("fma_fraction", C.c_double), ("sm_clock_mhz", C.c_double)]


class Template(C.Structure):
    _fields_ = [("prefix", C.c_uint8 * 136), ("prefix_len", C.c_uint32), ("prev_hash", C.c_uint8 * 32),
                ("n_int", C.c_int32), ("charset_count", C.c_int32)]


class Block(C.Structure):
    _fields_ = [("content", C.c_uint8 * 140), ("content_len", C.c_uint32), ("prev_hash", C.c_uint8 * 32),
                ("n_int", C.c_int32), ("charset_count", C.c_int32), ("reserved", C.c_uint32 * 2)]


class BatchItem(C.Structure):
    _fields_ = [("found", C.c_int32), ("nonce", C.c_uint32), ("digest", C.c_uint8 * 32)]


_lib = None


def library_path() -> str:
    return _build.LIB


def load():
    """dlopen the library (building it first if the sources are newer) and set the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if _build.needs_build():
        try:
            path = _build.build()
        except Exception as e:  # no nvcc on the GPU box is fine as long as a built .so travelled with the tree
            if not os.path.exists(path):
                raise UpowError(UPOW_E_INTERNAL, "load", f"libupow_b200.so missing and cannot be built: {e}")
    L = C.CDLL(path)
    vp, u8p, u32p, u64p = C.c_void_p, C.c_char_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
    L.upow_init.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(vp)]
    L.upow_destroy.argtypes = [vp]
    L.upow_device_count.argtypes = [vp]
    L.upow_get_device_info.argtypes = [vp, C.c_int, C.POINTER(DeviceInfo)]
    L.upow_last_error.argtypes = [vp]
    L.upow_last_error.restype = C.c_char_p
    L.upow_set_template.argtypes = [vp, C.c_int, u8p, C.c_int, u8p, C.c_int, C.c_int]
    L.upow_search_min.argtypes = [vp, C.c_int, C.c_uint64, C.c_uint64, C.POINTER(Limits), C.POINTER(Result)]
    L.upow_sweep.argtypes = [vp, C.c_int, C.c_uint64, C.c_uint64, u32p, C.c_uint64, u64p, C.POINTER(Result)]
    L.upow_search_batch.argtypes = [vp, C.c_uint32, C.POINTER(Template), C.c_uint64, C.POINTER(BatchItem),
                                    C.POINTER(Result)]
    L.upow_search_any.argtypes = [vp, C.c_uint32, C.POINTER(Template), C.c_uint64, C.POINTER(C.c_uint32),
                                  C.POINTER(BatchItem), C.POINTER(Result)]
    L.upow_verify_blocks.argtypes = [vp, C.c_uint32, C.POINTER(Block), C.c_char_p, C.c_char_p, C.POINTER(Result)]
    L.upow_sha256_batch.argtypes = [vp, C.c_uint32, C.c_char_p, u64p, C.c_char_p, C.POINTER(Result)]
    L.upow_verify_blocks_dev.argtypes = [vp, C.c_int, C.c_uint32, vp, vp, vp, C.POINTER(Result)]
    L.upow_sha256_batch_dev.argtypes = [vp, C.c_int, C.c_uint32, vp, vp, vp, C.POINTER(Result)]
    L.upow_cancel.argtypes = [vp]
    L.upow_clear_cancel.argtypes = [vp]
    L.upow_get_progress.argtypes = [vp, u64p]
    L.upow_selftest.argtypes = [C.c_int]
    L.upow_microbench_count.argtypes = []
    L.upow_microbench.argtypes = [C.c_int, C.c_int, C.c_uint32, C.POINTER(MicrobenchResult)]
    L.upow_flush_l2.argtypes = [vp]
    L.upow_version.restype = C.c_char_p
    L.upow_variant_name.argtypes = [C.c_int]
    L.upow_variant_name.restype = C.c_char_p
    L.upow_sweep_noalloc.argtypes = [C.c_int, u8p, C.c_int, u8p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, u32p,
                                     C.c_uint64, u64p]
    L.upow_noalloc_last_error.restype = C.c_char_p
    L.upow_get_variant.argtypes = [vp]
    L.upow_set_variant.argtypes = [vp, C.c_int]
    for name in EXPORTS:
        getattr(L, name)  # AttributeError here = header and library disagree
    _lib = L
    return L


class Context:
    """One upow_ctx: a set of GPUs plus template slots."""

    def __init__(self, n_devices: int = 1, device_ids=None):
        L = load()
        self._L = L
        self._h = C.c_void_p()
        ids = None
        if device_ids is not None:
            ids = (C.c_int * len(device_ids))(*device_ids)
            n_devices = len(device_ids)
        rc = L.upow_init(n_devices, ids, C.byref(self._h))
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_init", (L.upow_last_error(None) or b"").decode())

    # -- plumbing ---------------------------------------------------------------------
    def _err(self) -> str:
        return (self._L.upow_last_error(self._h) or b"").decode()

    def close(self):
        if getattr(self, "_h", None):
            self._L.upow_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_devices(self) -> int:
        return self._L.upow_device_count(self._h)

    def device_info(self, index: int = 0) -> DeviceInfo:
        info = DeviceInfo()
        rc = self._L.upow_get_device_info(self._h, index, C.byref(info))
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_get_device_info", self._err())
        return info

    # -- the path -----------------------------------------------------------------------
    def set_template(self, slot: int, prefix: bytes, prev_hash: bytes, n_int: int, charset_count: int):
        rc = self._L.upow_set_template(self._h, slot, bytes(prefix), len(prefix), bytes(prev_hash), n_int,
                                       charset_count)
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_set_template", self._err())

    def search_min(self, slot: int = 0, start_nonce: int = 0, stride: int = 1, max_nonce: int = 0,
                   max_rollovers: int = 0, time_limit_s: float = 0.0, tile_nonces: int = 0,
                   rollover_backwards: bool = False):
        """Returns (rc, Result); rc is UPOW_OK / UPOW_EXHAUSTED / UPOW_CANCELLED, errors raise."""
        lim = Limits(max_nonce, max_rollovers, 1 if rollover_backwards else 0, time_limit_s, tile_nonces)
        res = Result()
        rc = self._L.upow_search_min(self._h, slot, start_nonce, stride, C.byref(lim), C.byref(res))
        if rc < 0:
            raise UpowError(rc, "upow_search_min", self._err())
        return rc, res

    def sweep(self, slot: int, nonce_lo: int, nonce_hi: int, cap: int = 1 << 20):
        """Returns (sorted list of valid nonces, Result stats)."""
        buf = (C.c_uint32 * cap)()
        count = C.c_uint64(0)
        stats = Result()
        rc = self._L.upow_sweep(self._h, slot, nonce_lo, nonce_hi, buf, cap, C.byref(count), C.byref(stats))
        if rc != UPOW_OK:      # errors, and UPOW_CANCELLED: a partial set must not pass for the whole one
            raise UpowError(rc, "upow_sweep", self._err())
        return list(buf[: count.value]), stats

    def search_batch(self, templates, max_nonce: int = 0):
        """templates: iterable of (prefix, prev_hash, n_int, charset_count), or a ready (Template * n) array from
        template_array().  Returns (items, stats)."""
        arr = templates if isinstance(templates, C.Array) else self._template_array(list(templates))
        n = len(arr)
        items = (BatchItem * n)()
        stats = Result()
        rc = self._L.upow_search_batch(self._h, n, arr, max_nonce, items, C.byref(stats))
        if rc != UPOW_OK:      # errors, and UPOW_CANCELLED: the results of an interrupted batch mean nothing
            raise UpowError(rc, "upow_search_batch", self._err())
        return [(bool(it.found), it.nonce, bytes(it.digest)) for it in items], stats

    def template_array(self, templates):
        """The C array upow_search_batch / upow_search_any take, built once (4096 templates cost ~10 ms in Python)."""
        return self._template_array(list(templates))

    @staticmethod
    def _template_array(templates):
        arr = (Template * len(templates))()
        for t, (prefix, prev_hash, n_int, cc) in zip(arr, templates):
            C.memmove(t.prefix, bytes(prefix), len(prefix))
            t.prefix_len = len(prefix)
            C.memmove(t.prev_hash, bytes(prev_hash), 32)
            t.n_int, t.charset_count = n_int, cc
        return arr

    def search_any(self, templates, max_nonce: int = 0):
        """First success over many templates.  Returns (rc, index, nonce, digest, stats); rc 0 = found, 1 = none."""
        arr = templates if isinstance(templates, C.Array) else self._template_array(list(templates))
        idx, item, stats = C.c_uint32(0), BatchItem(), Result()
        rc = self._L.upow_search_any(self._h, len(arr), arr, max_nonce, C.byref(idx), C.byref(item),
                                     C.byref(stats))
        if rc < 0:
            raise UpowError(rc, "upow_search_any", self._err())
        return rc, idx.value, item.nonce, bytes(item.digest), stats

    def verify_blocks(self, blocks, want_digests: bool = True):
        """blocks: iterable of (content bytes, prev_hash bytes, n_int, charset_count).
        Returns (verdicts list of 1/0/255, digests list of bytes or None, stats)."""
        blocks = list(blocks)
        arr = (Block * max(len(blocks), 1))()
        for rec, (content, prev_hash, n_int, cc) in zip(arr, blocks):
            content = bytes(content)
            C.memmove(rec.content, content[:140], min(len(content), 140))
            rec.content_len = len(content)
            C.memmove(rec.prev_hash, bytes(prev_hash), 32)
            rec.n_int, rec.charset_count = n_int, cc
        valid = C.create_string_buffer(max(len(blocks), 1))
        dig = C.create_string_buffer(32 * max(len(blocks), 1)) if want_digests else None
        stats = Result()
        rc = self._L.upow_verify_blocks(self._h, len(blocks), arr, valid, dig, C.byref(stats))
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_verify_blocks", self._err())
        verdicts = list(valid.raw[: len(blocks)])
        digests = [dig.raw[32 * i: 32 * i + 32] for i in range(len(blocks))] if want_digests else None
        return verdicts, digests, stats

    def sha256_many(self, messages):
        """SHA-256 of every message (bytes) in one device call.  Returns (list of 32-byte digests, stats)."""
        messages = [bytes(m) for m in messages]
        n = len(messages)
        offs = (C.c_uint64 * (n + 1))()
        total = 0
        for i, m in enumerate(messages):
            offs[i] = total
            total += len(m)
        offs[n] = total
        out = C.create_string_buffer(32 * max(n, 1))
        stats = Result()
        rc = self._L.upow_sha256_batch(self._h, n, b"".join(messages), offs, out, C.byref(stats))
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_sha256_batch", self._err())
        return [out.raw[32 * i: 32 * i + 32] for i in range(n)], stats

    def cancel(self):
        """Sticky: a cancel issued while nothing runs makes the NEXT search return UPOW_CANCELLED."""
        self._L.upow_cancel(self._h)

    def clear_cancel(self):
        self._L.upow_clear_cancel(self._h)

    def progress(self) -> int:
        """Hashes evaluated so far by the running (or last) search_min; callable from another thread."""
        n = C.c_uint64(0)
        self._L.upow_get_progress(self._h, C.byref(n))
        return n.value

    @property
    def default_variant(self) -> int:
        return self._L.upow_default_variant()

    @property
    def variant(self) -> int:
        return self._L.upow_get_variant(self._h)

    @variant.setter
    def variant(self, v: int):
        rc = self._L.upow_set_variant(self._h, v)
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_set_variant")

    def variants(self):
        return [self._L.upow_variant_name(i).decode() for i in range(self._L.upow_variant_count())]

    def flush_l2(self):
        rc = self._L.upow_flush_l2(self._h)
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_flush_l2", self._err())


def sweep_noalloc(device: int, prefix: bytes, prev_hash: bytes, n_int: int, charset_count: int, nonce_lo: int,
                  nonce_hi: int, cap: int = NOALLOC_MAX_HITS):
    """upow_sweep_noalloc: sweep without any device allocation (for hosts that may not allocate; see the header).
    Returns the sorted valid nonces.  The output buffer here is ordinary host memory owned by ctypes."""
    L = load()
    buf = (C.c_uint32 * max(cap, 1))()
    count = C.c_uint64(0)
    rc = L.upow_sweep_noalloc(device, bytes(prefix), len(prefix), bytes(prev_hash), n_int, charset_count, nonce_lo,
                              nonce_hi, buf, cap, C.byref(count))
    if rc != UPOW_OK:
        raise UpowError(rc, "upow_sweep_noalloc", (L.upow_noalloc_last_error() or b"").decode())
    return list(buf[: count.value])


def selftest(device: int = 0) -> int:
    return load().upow_selftest(device)


def microbench(device: int = 0, iters: int = 2000):
    L = load()
    out = []
    for op in range(L.upow_microbench_count()):
        r = MicrobenchResult()
        rc = L.upow_microbench(device, op, iters, C.byref(r))
        if rc != UPOW_OK:
            raise UpowError(rc, "upow_microbench")
        out.append({"name": r.name.decode(), "warp_inst_per_clk_per_sm": r.warp_inst_per_clk_per_sm,
                    "lanes_per_clk_per_sm": r.lanes_per_clk_per_sm, "fma_fraction": r.fma_fraction,
                    "sm_clock_mhz": r.sm_clock_mhz, "kernel_ms": r.kernel_ms, "sm_count": r.sm_count})
    return out
