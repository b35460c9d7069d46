"""Host logic (no GPU): template preparation, the kernels' per-nonce arithmetic replayed on the
CPU through tests/host_harness.cpp, the Python template module, the C-ABI symbol table, and the
miner.py mirror's control flow with a stubbed context."""
import ctypes as C
import hashlib
import os
import random
import re
import sys

import pytest

from oracle import upow_oracle as O
from conftest import ROOT, has_gpu


class Tail(C.Structure):
    _fields_ = [("mid", C.c_uint32 * 8), ("w", C.c_uint32 * 16), ("pre", C.c_uint32 * 8), ("layout", C.c_uint32),
                ("pad", C.c_uint32)]


class Rule(C.Structure):
    _fields_ = [("want", C.c_uint32 * 8), ("n_target", C.c_uint32), ("frac_index", C.c_uint32),
                ("charset_count", C.c_uint32), ("sub0", C.c_uint32), ("range0", C.c_uint32)]


def _prepare(hh, prefix, prev_hash_hex, n_int, cc):
    t, r = Tail(), Rule()
    rc = hh.hh_prepare(prefix, len(prefix), bytes.fromhex(prev_hash_hex), n_int, cc, C.byref(t), C.byref(r))
    return rc, t, r


def test_struct_sizes(host_harness):
    a, b = C.c_int(), C.c_int()
    host_harness.hh_sizes(C.byref(a), C.byref(b))
    assert a.value == C.sizeof(Tail) and b.value == C.sizeof(Rule)


def test_host_sha256(host_harness):
    rng = random.Random(3)
    for n in (0, 3, 55, 56, 64, 108, 138, 777):
        m = rng.randbytes(n)
        out = C.create_string_buffer(32)
        host_harness.hh_sha256(m, n, out)
        assert out.raw == hashlib.sha256(m).digest()
        # the pipe-assigned compression of the verifier / batched-hashing kernels (portable fallbacks on the CPU)
        out2 = C.create_string_buffer(32)
        host_harness.hh_sha256_tuned(m, n, out2)
        assert out2.raw == hashlib.sha256(m).digest()


@pytest.mark.parametrize("version", [1, 2])
def test_midstate_tail_and_hot_path(host_harness, version):
    """Tail::mid/w/pre and the PRMT word builders give hashlib's digest word 0 for every nonce shape."""
    host_harness.hh_hot_a64.restype = C.c_uint32
    host_harness.hh_strided_a64.restype = C.c_uint32
    host_harness.hh_tuned_a64.restype = C.c_uint32
    rng = random.Random(10 + version)
    for trial in range(20):
        prefix = rng.randbytes(134 if version == 1 else 104)
        rc, t, r = _prepare(host_harness, prefix, O.T0_PREV_HASH, 5, 16)
        assert rc == 0 and t.layout == version
        py_mid, py_tail = O.midstate_and_tail(prefix)
        assert list(t.mid) == py_mid
        assert b"".join(x.to_bytes(4, "big") for x in t.w) == py_tail
        nonces = [0, 1, 0xff, 0x100, 0xffff, 0x10000, 0x00ff00ff, 0xff00ff00, 0x12345678, 0xffffffff] + \
                 [rng.getrandbits(32) for _ in range(20)]
        for nonce in nonces:
            want = hashlib.sha256(prefix + nonce.to_bytes(4, "little")).digest()
            d = (C.c_uint32 * 8)()
            host_harness.hh_full_digest(C.byref(t), nonce, d)
            assert b"".join(x.to_bytes(4, "big") for x in d) == want
            w0 = int.from_bytes(want[:4], "big")
            assert (host_harness.hh_hot_a64(C.byref(t), nonce) + t.mid[0]) & 0xffffffff == w0
            assert (host_harness.hh_strided_a64(C.byref(t), nonce) + t.mid[0]) & 0xffffffff == w0
            assert (host_harness.hh_tuned_a64(C.byref(t), nonce) + t.mid[0]) & 0xffffffff == w0


def test_rule_nibble_form_matches_reference_rule(host_harness):
    """rule_accepts == miner.py's string predicate for every N in 0..17 x every fraction, on digests
    engineered to be near the decision boundary; and the word-0 pre-filter never rejects a valid digest
    and is exact for N < 8."""
    rng = random.Random(5)
    prev = O.T0_PREV_HASH
    prefix = O.t0_prefix(4.0)
    for n in range(0, 18):
        for k in range(10):
            d = float(f"{n}.{k}")
            rule = O.Rule(d, prev)
            rc, t, r = _prepare(host_harness, prefix, prev, rule.int_difficulty, rule.charset_count)
            assert rc == 0
            assert r.n_target == len(rule.target_prefix)
            for trial in range(60):
                # start from the target, then damage 0-2 nibbles near the boundary
                hexd = list(rule.target_prefix + "".join(rng.choice("0123456789abcdef") for _ in range(64)))[:64]
                for _ in range(rng.randrange(0, 3)):
                    pos = min(63, max(0, len(rule.target_prefix) + rng.randrange(-3, 2)))
                    hexd[pos] = rng.choice("0123456789abcdef")
                hexd = "".join(hexd)
                words = (C.c_uint32 * 8)(*[int(hexd[8 * i:8 * i + 8], 16) for i in range(8)])
                want = rule.is_valid_hex(hexd)
                assert bool(host_harness.hh_rule_accepts(C.byref(r), words)) == want, (d, hexd)
                a64 = (words[0] - t.mid[0]) & 0xffffffff
                pre = bool(host_harness.hh_prefilter(C.byref(r), a64))
                if want:
                    assert pre
                if 1 <= n < 8:
                    assert pre == want, (d, hexd)


def test_prepare_rejects_bad_arguments(host_harness):
    p = O.t0_prefix(4.0)
    assert _prepare(host_harness, p[:-1], O.T0_PREV_HASH, 4, 16)[0] != 0   # 133 bytes
    assert _prepare(host_harness, p, O.T0_PREV_HASH, 4, 0)[0] != 0
    assert _prepare(host_harness, p, O.T0_PREV_HASH, 4, 17)[0] != 0
    assert _prepare(host_harness, p, O.T0_PREV_HASH, -1, 16)[0] != 0
    assert _prepare(host_harness, p, O.T0_PREV_HASH, 64, 8)[0] != 0    # digest_hex[64] would raise in the reference
    assert _prepare(host_harness, p, O.T0_PREV_HASH, 70, 16)[0] == 0   # s[-70:] is the whole string


def test_timestamp_helpers(host_harness):
    host_harness.hh_get_timestamp.restype = C.c_uint32
    for version in (1, 2):
        p = bytearray(O.t0_prefix(4.0, version))
        buf = (C.c_uint8 * len(p)).from_buffer(p)
        assert host_harness.hh_get_timestamp(buf, len(p)) == O.T0_TIMESTAMP
        host_harness.hh_set_timestamp(buf, len(p), O.T0_TIMESTAMP + 5)
        assert bytes(p) == O.set_timestamp(O.t0_prefix(4.0, version), O.T0_TIMESTAMP + 5)


def test_python_template_module_matches_oracle(mine_block_vectors):
    from upow_b200 import template as T
    for v in mine_block_vectors:
        p = v["mining_params"]
        prev = p["last_block"].get("hash", T.GENESIS_PREV_HASH)
        rule = T.DifficultyRule.derive(p["difficulty"], prev)
        ref = O.Rule(p["difficulty"], prev)
        assert (rule.n_int, rule.charset_count, rule.target_prefix) == \
               (ref.int_difficulty, ref.charset_count, ref.target_prefix)
        prefix = T.build_block_prefix(prev, T.string_to_bytes(v["wallet"]),
                                      T.calculate_merkle_root(p["pending_transactions_hashes"]), v["timestamp"],
                                      p["difficulty"])
        assert prefix.hex() == v["block_content"][:-8]
        assert rule.accepts(v["digest"])
        assert T.with_timestamp(prefix, v["timestamp"] + 3) == O.set_timestamp(prefix, v["timestamp"] + 3)


def test_library_exports_every_declared_symbol():
    import upow_b200
    header = open(os.path.join(ROOT, "include", "upow_b200.h")).read()
    declared = set(re.findall(r"\b(upow_[a-z0-9_]+)\s*\(", header))
    assert {"upow_init", "upow_set_template", "upow_search_min", "upow_sweep", "upow_search_batch", "upow_cancel",
            "upow_selftest", "upow_destroy", "upow_last_error"} <= declared
    lib = upow_b200.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/upow_b200.h but not exported"
    assert b"sm_100a" in lib.upow_version()


def test_ctypes_struct_layouts_match_header():
    """Field-by-field size check of the ctypes mirrors against a C compile of the header."""
    import subprocess, tempfile
    from upow_b200 import _lib
    src = '#include <stdio.h>\n#include "upow_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",' \
          'sizeof(upow_limits),sizeof(upow_result),sizeof(upow_device_info),sizeof(upow_microbench_result),' \
          'sizeof(upow_template),sizeof(upow_batch_item),__builtin_offsetof(upow_result,hashes_done));return 0;}'
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "s.c")
        open(c, "w").write(src)
        exe = os.path.join(td, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", exe, c])
        sizes = list(map(int, subprocess.check_output([exe]).split()))
    assert sizes == [C.sizeof(_lib.Limits), C.sizeof(_lib.Result), C.sizeof(_lib.DeviceInfo),
                     C.sizeof(_lib.MicrobenchResult), C.sizeof(_lib.Template), C.sizeof(_lib.BatchItem),
                     _lib.Result.hashes_done.offset]


@pytest.mark.skipif(has_gpu(), reason="only meaningful on a machine without a GPU")
def test_no_cpu_fallback_without_gpu():
    import upow_b200
    with pytest.raises(upow_b200.UpowError) as e:
        upow_b200.Context(1)
    assert e.value.code == -2


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "upow_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+\.*oracle", text, re.M), f
                assert "libupow_oracle" not in text and "upow_oracle" not in text and "oracle/" not in text, f


class _FakeResult:
    def __init__(self, nonce, ts, hashes=1000):
        self.nonce, self.timestamp, self.hashes_done, self.found = nonce, ts, hashes, 1


class _FakeContext:
    """Stands in for upow_b200.Context: answers from the oracle.  Only exercises miner.py's control flow."""

    def __init__(self):
        self.calls = []

    def set_template(self, slot, prefix, prev_hash, n_int, cc):
        self.prefix, self.prev_hash, self.n_int, self.cc = prefix, prev_hash, n_int, cc

    def search_min(self, slot, start, stride, time_limit_s=0, max_rollovers=0, **kw):
        self.calls.append((start, stride, time_limit_s))
        tgt = self.prev_hash.hex()[-self.n_int:]
        n = start
        while n < 2 ** 32:
            h = hashlib.sha256(self.prefix + n.to_bytes(4, "little")).hexdigest()
            if h.startswith(tgt) and (self.cc == 16 or int(h[self.n_int], 16) < self.cc):
                return 0, _FakeResult(n, O.get_timestamp(self.prefix))
            n += stride
        return 1, _FakeResult(0, 0)


def test_miner_mirror_call_surface(monkeypatch, capsys, mine_block_vectors):
    """mine_block(start, inc, params) posts the same JSON the reference posts (golden vectors)."""
    import types
    from upow_b200 import miner
    posted = {}

    class Resp:
        def json(self):
            return {"ok": True}

    fake_requests = types.ModuleType("requests")
    fake_requests.post = lambda url, json=None, timeout=None: (posted.update(json, url=url, timeout=timeout), Resp())[1]
    monkeypatch.setitem(sys.modules, "requests", fake_requests)
    for v in [x for x in mine_block_vectors if x["nonce"] < 700_000][:10]:
        posted.clear()
        ctx = _FakeContext()
        miner.set_context(ctx)
        monkeypatch.setattr(sys, "argv", ["miner.py", v["wallet"], "1", "http://node.example:3006"])
        monkeypatch.setattr(miner, "timestamp", lambda ts=v["timestamp"]: ts)
        params = {k: (dict(val) if isinstance(val, dict) else val) for k, val in v["mining_params"].items()}
        with pytest.raises(SystemExit):
            miner.mine_block(v["start_nonce"], v["nonce_increment"], params)
        assert posted["block_content"] == v["block_content"], v["name"]
        assert posted["txs"] == v["posted_txs"] and posted["block_no"] == v["posted_block_no"]
        assert posted["url"] == "http://node.example:3006/push_block"
        assert posted["timeout"] == 20 + int((len(v["posted_txs"]) or 1) / 3)
        assert ctx.calls == [(v["start_nonce"], v["nonce_increment"], 90)]
        out = capsys.readouterr().out
        assert v["block_content"] in out and "BLOCK MINED" in out
        if v["start_nonce"] == 0:
            assert f"difficulty: {v['mining_params']['difficulty']}" in out
    miner.set_context(None)


def test_miner_mirror_exhaustion_raises_overflow(monkeypatch):
    import types
    from upow_b200 import miner, _lib

    class Ctx(_FakeContext):
        def search_min(self, *a, **k):
            return _lib.UPOW_EXHAUSTED, _FakeResult(0, 0)

    monkeypatch.setitem(sys.modules, "requests", types.ModuleType("requests"))
    miner.set_context(Ctx())
    monkeypatch.setattr(sys, "argv", ["miner.py", O.T0_ADDRESS_V1])
    with pytest.raises(OverflowError):  # the reference: nonce.to_bytes(4, ...) at 2**32, miner.py:87
        miner.mine_block(0, 1, O.t0_mining_params(4.0))
    miner.set_context(None)


def test_committed_sass_counts_match_the_built_library():
    """profiles/sass_counts.json (what bench.py's roofline uses) is in step with libupow_b200.so."""
    import json
    import shutil
    import upow_b200
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not installed")
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import sass_count
    committed = json.load(open(os.path.join(ROOT, "profiles", "sass_counts.json")))
    name = committed["bench_kernel"]
    assert name and name in committed["kernels"]
    upow_b200.load()
    live = sass_count.analyse(upow_b200.library_path(), r"search_kernel_tunedILi1ELi1ENS_4TuneILi1ELi0ELi0ELi0ELi0ELi0ELi0ELi4EEELi3ELi256")
    assert len(live) == 1
    (r,) = live.values()
    k = committed["kernels"][name]
    assert (r["total"], r["pipes"].get("alu", 0), r["pipes"].get("fma", 0)) == \
           (k["total_per_hash"], k["alu_per_hash"], k["fma_per_hash"])
    # the generic model of SURVEY 8(d): between the folded estimate -10 % and 1386 + pipe-move overhead
    assert 1100 < k["total_per_hash"] < 1600 and k["alu_per_hash"] < 1000
    # external anchor of the roofline: the shipped kernels stay within 2 % of the algorithmic
    # floor of alu-only instructions (rotates, shifts, three-input logic), for both layouts and both modes
    import bench
    for layout, floor in bench.ALU_FLOOR_PER_HASH.items():
        for mode in (0, 1):
            live = sass_count.analyse(upow_b200.library_path(),
                                      rf"search_kernel_tunedILi{layout}ELi{mode}ENS_4TuneILi1ELi0ELi0ELi0ELi0ELi0ELi0ELi4EEELi3ELi256")
            (r,) = live.values()
            assert floor <= r["pipes"]["alu"] <= floor * 1.02, (layout, mode, r["pipes"])


def test_miner_main_round_structure_with_multiplier(monkeypatch, capsys):
    """main(): fetch get_mining_info, race UPOW_TEMPLATES templates, POST the winner, start the next round
    (miner.py:126-156 with the optional template multiplier)."""
    import types
    from upow_b200 import miner

    class AnyCtx(_FakeContext):
        def search_any(self, templates, max_nonce=0):
            for n in range(1 << 20):
                for i, (prefix, prev, n_int, cc) in enumerate(templates):
                    h = hashlib.sha256(prefix + n.to_bytes(4, "little")).hexdigest()
                    if h.startswith(prev.hex()[-n_int:]) and (cc == 16 or int(h[n_int], 16) < cc):
                        class S:
                            hashes_done, elapsed_s = 1, 0.0
                        return 0, i, n, bytes.fromhex(h), S()
            return 1, 0, 0, b"", None

    info = {"difficulty": 2.5, "last_block": {"hash": O.T0_PREV_HASH, "id": 41, "timestamp": 1_699_999_990},
            "pending_transactions_hashes": sorted(hashlib.sha256(b"t%d" % i).hexdigest() for i in range(2))}
    posted, gets = [], []

    class Resp:
        def __init__(self, payload):
            self.payload = payload

        def json(self):
            return self.payload

    def fake_get(url, timeout=None):
        gets.append(url)
        if len(gets) > 1:
            raise KeyboardInterrupt          # end the endless round loop of main()
        return Resp({"ok": True, "result": info})

    fake = types.ModuleType("requests")
    fake.get = fake_get
    fake.post = lambda url, json=None, timeout=None: (posted.append((url, json)), Resp({"ok": True}))[1]
    monkeypatch.setitem(sys.modules, "requests", fake)
    monkeypatch.setenv("UPOW_TEMPLATES", "12")
    monkeypatch.setattr(sys, "argv", ["miner.py", O.T0_ADDRESS_V1, "1", "http://node.example:3006/"])
    monkeypatch.setattr(miner, "timestamp", lambda: 1_700_000_000)
    miner.set_context(AnyCtx())
    with pytest.raises(KeyboardInterrupt):
        miner.main()
    miner.set_context(None)
    assert gets[0] == "http://node.example:3006/get_mining_info" and len(posted) == 1
    url, body = posted[0]
    assert url.endswith("/push_block") and body["block_no"] == 42
    assert O.node_check_block_is_valid(body["block_content"], 2.5, info["last_block"])
    raw = bytes.fromhex(body["block_content"])
    assert 1_699_999_990 < int.from_bytes(raw[-10:-6], "little") <= 1_700_000_000
    assert raw[96:128].hex() == O.calculate_merkle_root(body["txs"])
    assert "BLOCK MINED" in capsys.readouterr().out


def test_null_arguments_are_rejected_without_a_gpu():
    """Every entry point checks its pointers before touching CUDA: NULL context / outputs give UPOW_E_ARG (-1)."""
    import upow_b200
    from upow_b200 import _lib
    L = upow_b200.load()
    res, lim = _lib.Result(), _lib.Limits()
    cnt = C.c_uint64(0)
    idx = C.c_uint32(0)
    item = _lib.BatchItem()
    assert L.upow_search_min(None, 0, 0, 1, C.byref(lim), C.byref(res)) == -1
    assert L.upow_sweep(None, 0, 0, 16, None, 0, C.byref(cnt), None) == -1
    assert L.upow_set_template(None, 0, b"x" * 134, 134, bytes(32), 4, 16) == -1
    assert L.upow_search_batch(None, 0, None, 0, None, None) == -1
    assert L.upow_search_any(None, 0, None, 0, C.byref(idx), C.byref(item), None) == -1
    assert L.upow_verify_blocks(None, 0, None, None, None, None) == -1
    assert L.upow_sha256_batch(None, 0, None, None, None, None) == -1
    assert L.upow_cancel(None) == -1 and L.upow_destroy(None) == -1 and L.upow_flush_l2(None) == -1
    assert L.upow_device_count(None) == -1 and L.upow_get_variant(None) == -1 and L.upow_set_variant(None, 0) == -1
    assert L.upow_init(99, None, C.byref(C.c_void_p())) == -1 and L.upow_init(1, None, None) == -1
    assert L.upow_microbench(0, -1, 10, None) == -1
    assert L.upow_sweep_noalloc(0, None, 134, bytes(32), 4, 16, 0, 16, None, 0, C.byref(cnt)) == -1
    assert L.upow_variant_count() >= 2 and b"plain" in L.upow_variant_name(0) and L.upow_variant_name(999) == b""
    assert 0 < L.upow_default_variant() < L.upow_variant_count()
    assert L.upow_clear_cancel(None) == -1
    assert L.upow_verify_blocks_dev(None, 0, 0, None, None, None, None) == -1
    assert L.upow_sha256_batch_dev(None, 0, 0, None, None, None, None) == -1


def test_miner_never_searches_an_exhausted_space_twice(monkeypatch, capsys):
    """A B200 exhausts the 2^32 nonces of one timestamp in 0.23 s; the reference's restart-with-a-fresh-timestamp
    (miner.py:87,149-156) would rebuild the SAME template inside the same second.  The mirror rolls the timestamp
    backwards through the node's window (upow/manager.py:445-460) in one call and remembers what it exhausted."""
    import types
    from upow_b200 import miner, _lib

    class ExhaustedCtx:
        def __init__(self):
            self.calls = []

        def set_template(self, *a):
            pass

        def search_min(self, slot, start, stride, **kw):
            self.calls.append(kw)
            return _lib.UPOW_EXHAUSTED, _FakeResult(0, 0)

    ctx = ExhaustedCtx()
    miner.set_context(ctx)
    miner._searched.clear()
    monkeypatch.setitem(sys.modules, "requests", types.ModuleType("requests"))
    monkeypatch.setattr(sys, "argv", ["miner.py", O.T0_ADDRESS_V1])
    T = 1_700_000_000
    params = {"difficulty": 9.0, "last_block": {"hash": O.T0_PREV_HASH, "id": 7, "timestamp": T - 20},
              "pending_transactions_hashes": []}
    monkeypatch.setattr(miner, "timestamp", lambda: T)
    with pytest.raises(OverflowError):
        miner.mine_block(0, 1, dict(params, last_block=dict(params["last_block"])))
    # one call covered every acceptable timestamp: T, T-1, ..., T-19 (20 spaces), newest first
    assert len(ctx.calls) == 1 and ctx.calls[0]["max_rollovers"] == 19 and ctx.calls[0]["rollover_backwards"] is True
    with pytest.raises(OverflowError):      # same second again: nothing new to search, the GPU is not asked
        miner.mine_block(0, 1, dict(params, last_block=dict(params["last_block"])))
    assert len(ctx.calls) == 1
    monkeypatch.setattr(miner, "timestamp", lambda: T + 3)
    with pytest.raises(OverflowError):      # three new seconds: exactly those are searched
        miner.mine_block(0, 1, dict(params, last_block=dict(params["last_block"])))
    assert len(ctx.calls) == 2 and ctx.calls[1]["max_rollovers"] == 2
    assert list(miner._searched.values()) == [(T + 3, T - 19)]
    # a new chain tip is a new search space
    params2 = dict(params, last_block={"hash": "cd" * 32, "id": 8, "timestamp": T + 2})
    with pytest.raises(OverflowError):
        miner.mine_block(0, 1, params2)
    assert len(ctx.calls) == 3 and ctx.calls[2]["max_rollovers"] == 0      # only T+3 lies after the new tip's timestamp
    miner.set_context(None)
    miner._searched.clear()
    capsys.readouterr()


def test_miner_main_throttles_get_mining_info(monkeypatch, capsys):
    """get_mining_info is rate limited to 30 a minute (upow/node/main.py:675-676): main() leaves at least
    MINING_INFO_MIN_PERIOD_S between two requests even when every round ends at once."""
    import types
    from upow_b200 import miner, _lib

    class Ctx:
        cleared = 0

        def set_template(self, *a):
            pass

        def search_min(self, *a, **k):
            return _lib.UPOW_EXHAUSTED, _FakeResult(0, 0)

        def clear_cancel(self):
            Ctx.cleared += 1

    clock = {"t": 1000.0}
    sleeps, gets = [], []
    monkeypatch.setattr(miner.time, "time", lambda: clock["t"])

    def fake_sleep(s):
        sleeps.append(s)
        clock["t"] += s
    monkeypatch.setattr(miner.time, "sleep", fake_sleep)

    class Resp:
        def json(self):
            return {"ok": True, "result": {"difficulty": 9.0, "last_block": {"hash": O.T0_PREV_HASH, "id": 1, "timestamp": 1_699_999_999},
                                           "pending_transactions_hashes": []}}

    def fake_get(url, timeout=None):
        gets.append(clock["t"])
        if len(gets) > 4:
            raise KeyboardInterrupt
        return Resp()

    fake = types.ModuleType("requests")
    fake.get = fake_get
    monkeypatch.setitem(sys.modules, "requests", fake)
    monkeypatch.setattr(sys, "argv", ["miner.py", O.T0_ADDRESS_V1, "1"])
    monkeypatch.setattr(miner, "timestamp", lambda: 1_700_000_000)
    monkeypatch.delenv("UPOW_TEMPLATES", raising=False)
    miner._searched.clear()
    miner.set_context(Ctx())
    with pytest.raises(KeyboardInterrupt):
        miner.main()
    miner.set_context(None)
    miner._searched.clear()
    assert all(b - a >= miner.MINING_INFO_MIN_PERIOD_S - 1e-9 for a, b in zip(gets, gets[1:]))
    assert Ctx.cleared == 4          # every fresh template drops a cancel left over from the previous one
    capsys.readouterr()
