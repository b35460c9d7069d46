"""Property tests (hypothesis, seeds derandomised so the runs are reproducible): the kernels' per-nonce
arithmetic on the CPU (host harness) and the CUDA search on windows small enough for the C oracle."""
import ctypes as C
import hashlib

import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from oracle import c_oracle
from oracle import upow_oracle as O
from test_host_logic import Rule, Tail

DIFFICULTIES = st.sampled_from([1.0, 1.3, 2.0, 2.5, 2.9, 3.0, 3.6, 4.0, 4.4, 5.0, 5.1, 6.0, 7.7, 8.0, 9.2, 12.5, 16.0])
PREFIX = st.one_of(st.binary(min_size=134, max_size=134), st.binary(min_size=104, max_size=104))
COMMON = dict(deadline=None, derandomize=True, suppress_health_check=[HealthCheck.function_scoped_fixture])


@settings(max_examples=150, **COMMON)
@given(prefix=PREFIX, prev=st.binary(min_size=32, max_size=32), nonce=st.integers(0, 2 ** 32 - 1), d=DIFFICULTIES)
def test_per_nonce_arithmetic_and_rule_on_cpu(host_harness, prefix, prev, nonce, d):
    """Every code path that produces digest word 0 (slow path, plain hot path, strided path, hand-specialised
    loops) equals hashlib; the pre-filter never rejects what the reference rule accepts."""
    for fn in ("hh_hot_a64", "hh_strided_a64", "hh_tuned_a64"):
        getattr(host_harness, fn).restype = C.c_uint32
    rule = O.Rule(d, prev.hex())
    t, r = Tail(), Rule()
    assert host_harness.hh_prepare(prefix, len(prefix), prev, rule.int_difficulty, rule.charset_count,
                                   C.byref(t), C.byref(r)) == 0
    digest = hashlib.sha256(prefix + nonce.to_bytes(4, "little")).digest()
    words = (C.c_uint32 * 8)()
    host_harness.hh_full_digest(C.byref(t), nonce, words)
    assert b"".join(w.to_bytes(4, "big") for w in words) == digest
    w0 = int.from_bytes(digest[:4], "big")
    for fn in ("hh_hot_a64", "hh_strided_a64", "hh_tuned_a64"):
        assert (getattr(host_harness, fn)(C.byref(t), nonce) + t.mid[0]) & 0xFFFFFFFF == w0
    valid = rule.is_valid_hex(digest.hex())
    assert bool(host_harness.hh_rule_accepts(C.byref(r), words)) == valid
    if valid:
        assert host_harness.hh_prefilter(C.byref(r), (w0 - t.mid[0]) & 0xFFFFFFFF)


@pytest.mark.gpu
@settings(max_examples=40, **COMMON)
@given(prefix=PREFIX, prev=st.binary(min_size=32, max_size=32), d=st.sampled_from([1.0, 1.6, 2.0, 2.7, 3.0, 3.3, 4.0]),
       lo=st.integers(0, 2 ** 32 - 1), span=st.integers(0, 1 << 18), stride=st.sampled_from([1, 1, 1, 2, 3, 7, 255]))
def test_windowed_search_and_sweep_match_the_c_oracle(gpu_ctx, prefix, prev, d, lo, span, stride):
    """Arbitrary windows [lo, lo+span) anywhere in the nonce space, both layouts, strides: same minimum (or
    'exhausted') as the C oracle; the sweep returns exactly the oracle's set."""
    hi = min(lo + span, 1 << 32)
    rule = O.Rule(d, prev.hex())
    gpu_ctx.set_template(3, prefix, prev, rule.int_difficulty, rule.charset_count)
    want, _ = c_oracle.min_valid_nonce(prefix, rule, lo, stride, hi)
    rc, res = gpu_ctx.search_min(3, lo, stride, max_nonce=hi) if hi > 0 else (1, None)
    if want is None:
        assert rc == 1
    else:
        assert rc == 0 and res.nonce == want
        assert bytes(res.digest) == hashlib.sha256(prefix + want.to_bytes(4, "little")).digest()
    got, _ = gpu_ctx.sweep(3, lo, hi)
    assert got == c_oracle.all_valid_nonces(prefix, rule, lo, hi)[0]


@pytest.mark.gpu
@settings(max_examples=25, **COMMON)
@given(prefix=PREFIX, prev=st.binary(min_size=32, max_size=32), d=st.sampled_from([2.0, 2.5, 3.0, 3.4]),
       cap=st.integers(64, 3000), backwards=st.booleans(), tile=st.sampled_from([0, 1 << 16]))
def test_rollover_search_matches_the_oracle_in_both_directions(gpu_ctx, prefix, prev, d, cap, backwards, tile):
    """Timestamp rollover over a nonce space cut to `cap`: the first valid (rollover index, nonce) in search order,
    forwards (benchmark order) and backwards (what a live miner uses), against the hashlib loop per timestamp."""
    rule = O.Rule(d, prev.hex())
    ts0 = O.get_timestamp(prefix)
    max_roll = 40
    want = None
    for k in range(max_roll + 1):
        ts = (ts0 - k if backwards else ts0 + k) & 0xFFFFFFFF
        n = O.min_valid_nonce(O.set_timestamp(prefix, ts), rule, 0, 1, cap)
        if n is not None:
            want = (k, ts, n)
            break
    gpu_ctx.set_template(4, prefix, prev, rule.int_difficulty, rule.charset_count)
    rc, res = gpu_ctx.search_min(4, max_nonce=cap, max_rollovers=max_roll, rollover_backwards=backwards, tile_nonces=tile)
    if want is None:
        assert rc == 1 and res.found == 0
    else:
        assert rc == 0 and (res.rollovers, res.timestamp, res.nonce) == want
        assert bytes(res.digest) == O.digest_of(O.set_timestamp(prefix, want[1]), want[2])


@pytest.mark.gpu
@settings(max_examples=12, **COMMON)
@given(seeds=st.lists(st.integers(0, 2 ** 31), min_size=1, max_size=40, unique=True),
       d=st.sampled_from([1.5, 2.0, 3.0, 3.7]), cap_log2=st.sampled_from([12, 16, 32]))
def test_batch_engine_matches_the_c_oracle_on_random_templates(gpu_ctx, seeds, d, cap_log2):
    """Random templates of both layouts in one batch, bounded and unbounded nonce space: every minimum (or
    'none below the bound') equals the C oracle's; search_any picks the lowest index of the earliest round."""
    import random
    temps, wants = [], []
    for s in seeds:
        rng = random.Random(s)
        prefix = rng.randbytes(134 if s % 3 else 104)
        prev = rng.randbytes(32)
        rule = O.Rule(d, prev.hex())
        temps.append((prefix, prev, rule.int_difficulty, rule.charset_count))
        wants.append(c_oracle.min_valid_nonce(prefix, rule, 0, 1, 1 << cap_log2, threads=1)[0])
    items, _ = gpu_ctx.search_batch(temps, max_nonce=(1 << cap_log2) & 0xFFFFFFFF if cap_log2 < 32 else 0)
    for (found, nonce, digest), want, t in zip(items, wants, temps):
        assert (found, nonce if found else None) == (want is not None, want)
        if found:
            assert digest == hashlib.sha256(t[0] + nonce.to_bytes(4, "little")).digest()
    rc, idx, nonce, digest, _ = gpu_ctx.search_any(temps, max_nonce=(1 << cap_log2) if cap_log2 < 32 else 0)
    if all(w is None for w in wants):
        assert rc == 1
    else:
        assert rc == 0 and nonce == wants[idx]
