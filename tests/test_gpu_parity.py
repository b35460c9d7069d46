"""Parity of the CUDA path with the oracle, through the C-ABI (ctypes -> libupow_b200.so).
Bit-exact everywhere: nonce, timestamp, digest, valid-nonce sets."""
import hashlib
import random
import threading
import time

import pytest

from oracle import c_oracle
from oracle import upow_oracle as O

pytestmark = pytest.mark.gpu


class _Rule(O.Rule):
    """oracle rule that also remembers the previous hash it was derived from"""
    __slots__ = ("prev",)


def make_rule(difficulty, prev):
    r = _Rule(difficulty, prev)
    r.prev = prev
    return r


def _set(ctx, slot, prefix, rule):
    ctx.set_template(slot, prefix, bytes.fromhex(rule.prev), rule.int_difficulty, rule.charset_count)


def test_selftest_kats():
    import upow_b200
    assert upow_b200.selftest(0) == 0


def test_device_info(gpu_ctx):
    info = gpu_ctx.device_info(0)
    assert info.sm_count > 0 and info.search_grid_blocks % info.sm_count == 0
    assert info.cc_major >= 10, "this library is built for sm_100a only"


def test_golden_mine_block_vectors(gpu_ctx, mine_block_vectors):
    """configs[0], configs[1] and all other reference-generated vectors: same nonce, same digest, same block."""
    for v in mine_block_vectors:
        p = v["mining_params"]
        prev = p["last_block"].get("hash", O.GENESIS_PREV_HASH)
        rule = make_rule(p["difficulty"], prev)
        prefix = bytes.fromhex(v["block_content"])[:-4]
        _set(gpu_ctx, 0, prefix, rule)
        rc, res = gpu_ctx.search_min(0, v["start_nonce"], v["nonce_increment"])
        assert rc == 0 and res.found == 1, v["name"]
        assert res.nonce == v["nonce"], v["name"]
        assert res.digest_hex == v["digest"], v["name"]
        assert res.timestamp == v["timestamp"] and res.rollovers == 0
        assert (prefix + res.nonce.to_bytes(4, "little")).hex() == v["block_content"]
        assert O.node_check_block_is_valid(v["block_content"], p["difficulty"], {"hash": prev})
        assert res.hashes_done > 0 and res.launches >= 1


def test_config2_fractional_6_3(gpu_ctx):
    prefix = O.t0_prefix(6.3)
    _set(gpu_ctx, 0, prefix, make_rule(6.3, O.T0_PREV_HASH))
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 9436117)
    assert res.digest_hex == "3745dc106adcf74ca0691a8f6031a651ef63add4c44d78962e83dfc36c884343"


@pytest.mark.parametrize("version", [1, 2])
def test_random_templates_min_nonce(gpu_ctx, version):
    rng = random.Random(100 + version)
    for i in range(24):
        d = rng.choice([1.0, 2.4, 3.0, 3.9, 4.0, 4.1, 4.7, 5.0, 5.5, 5.9, 6.0])
        prev = rng.randbytes(32).hex()
        prefix = rng.randbytes(134 if version == 1 else 104)
        rule = make_rule(d, prev)
        _set(gpu_ctx, 1, prefix, rule)
        rc, res = gpu_ctx.search_min(1)
        want, _ = c_oracle.min_valid_nonce(prefix, rule)
        if want is None:
            assert rc == 1 and res.found == 0
            continue
        assert rc == 0 and res.nonce == want, (version, i, d)
        assert bytes(res.digest) == hashlib.sha256(prefix + want.to_bytes(4, "little")).digest()


@pytest.mark.parametrize("version", [1, 2])
def test_strided_search_matches_reference_worker_layout(gpu_ctx, version):
    """(start, inc) = (k, W): minimum over the residue class, as worker k of W finds it (miner.py:139-148)."""
    rng = random.Random(200 + version)
    prefix = O.t0_prefix(4.6, version)
    rule = make_rule(4.6, O.T0_PREV_HASH)
    _set(gpu_ctx, 0, prefix, rule)
    default_variant = gpu_ctx.variant
    try:
        for variant in (default_variant, 0):      # pipe-assigned build and the plain-CUDA build of the strided kernel
            gpu_ctx.variant = variant
            for start, inc in [(0, 2), (1, 2), (3, 8), (7, 8), (5, 96), (12345, 1000), (0, 65536), (17, 65537), (2, 3)]:
                rc, res = gpu_ctx.search_min(0, start, inc)
                want, _ = c_oracle.min_valid_nonce(prefix, rule, start, inc)
                assert (rc == 0 and res.nonce == want) if want is not None else rc == 1, (variant, start, inc)
                if want is not None:
                    assert (res.nonce - start) % inc == 0
    finally:
        gpu_ctx.variant = default_variant


def test_start_nonce_and_max_nonce_bounds(gpu_ctx, sweep_fixture):
    """Contiguous search from an arbitrary start / up to an arbitrary end: edges are exact."""
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    _set(gpu_ctx, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
    ns = sweep_fixture["nonces"]
    for idx in (0, 1, 5, 100, 2000, 4208):
        n = ns[idx]
        rc, res = gpu_ctx.search_min(0, n)          # start exactly on a valid nonce
        assert (rc, res.nonce) == (0, n)
        rc, res = gpu_ctx.search_min(0, n + 1)      # just after it
        if idx + 1 < len(ns):
            assert (rc, res.nonce) == (0, ns[idx + 1])
        else:
            assert rc == 1 and res.found == 0       # range exhausted (reference: OverflowError)
        rc, res = gpu_ctx.search_min(0, ns[idx - 1] + 1 if idx else 0, max_nonce=n)   # end just before it
        assert rc == 1 and res.found == 0
        rc, res = gpu_ctx.search_min(0, ns[idx - 1] + 1 if idx else 0, max_nonce=n + 1)
        assert (rc, res.nonce) == (0, n)
    rc, res = gpu_ctx.search_min(0, 2 ** 32)  # empty range
    assert rc == 1 and res.hashes_done == 0


def test_sweep_slices_match_oracle(gpu_ctx, sweep_fixture):
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    _set(gpu_ctx, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
    ns = sweep_fixture["nonces"]
    for lo, hi in [(0, 0), (0, 1), (164867, 164868), (164868, 2752366), (0, 1 << 16), (12345, 7_000_001),
                   (1 << 28, (1 << 28) + (1 << 24) + 17), ((1 << 32) - 5_000_000, 1 << 32)]:
        got, stats = gpu_ctx.sweep(0, lo, hi)
        assert got == [n for n in ns if lo <= n < hi], (lo, hi)


def test_config3_full_sweep_exact_set(gpu_ctx, sweep_fixture):
    """Full 2^32 sweep at difficulty 5.0: exactly the 4209 nonces hashlib finds."""
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    _set(gpu_ctx, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
    got, stats = gpu_ctx.sweep(0, 0, 1 << 32)
    assert stats.hashes_done == 1 << 32
    assert len(got) == 4209
    assert got == sweep_fixture["nonces"]
    assert O.nonce_set_hash(got) == "5ca83877d72644292a359a6a2864d9411523a9c01a2f2a0ec6aea1ee3c9d742f"


def test_sweep_v2_and_fractional_against_c_oracle(gpu_ctx):
    for version, d in ((2, 4.0), (1, 3.5), (2, 4.7), (1, 2.0)):
        prefix = O.t0_prefix(d, version)
        rule = make_rule(d, O.T0_PREV_HASH)
        _set(gpu_ctx, 2, prefix, rule)
        lo, hi = 1_000_003, 1_000_003 + (1 << 23)
        got, _ = gpu_ctx.sweep(2, lo, hi)
        want, _ = c_oracle.all_valid_nonces(prefix, rule, lo, hi)
        assert got == want and len(want) > 0, (version, d)


def test_sweep_overflow_is_an_error(gpu_ctx):
    import upow_b200
    prefix = O.t0_prefix(1.0)
    _set(gpu_ctx, 0, prefix, make_rule(1.0, O.T0_PREV_HASH))
    with pytest.raises(upow_b200.UpowError) as e:
        gpu_ctx.sweep(0, 0, 1 << 20, cap=16)
    assert e.value.code == -3


def test_high_difficulty_rules_word1_and_quirks(gpu_ctx):
    """N >= 8 (second digest word), N == 0 quirk (whole previous hash), int-valued difficulty."""
    prefix = O.t0_prefix(4.0)
    n0 = 62804
    digest = O.digest_of(prefix, n0).hex()
    for n_int, cc_nib in ((8, 16), (9, 16), (12, 16), (16, 16), (17, 16), (10, None), (0, 16)):
        # craft a prev hash whose tail equals the digest head, so n0 is valid at this difficulty
        if n_int == 0:
            prev = digest
        else:
            prev = "ab" * 32
            prev = prev[:64 - n_int] + digest[:n_int]
        cc = 16 if cc_nib else int(digest[n_int], 16) + 1
        gpu_ctx.set_template(0, prefix, bytes.fromhex(prev), n_int, cc)
        rc, res = gpu_ctx.search_min(0, 0, 1, max_nonce=1 << 20)
        assert (rc, res.nonce) == (0, n0), n_int
        if cc < 16 and cc > 1:
            gpu_ctx.set_template(0, prefix, bytes.fromhex(prev), n_int, cc - 1)   # nibble now just outside the charset
            rc, res = gpu_ctx.search_min(0, 0, 1, max_nonce=1 << 20)
            assert rc == 1
        # one nibble off in the deepest constrained position: no longer valid
        if n_int:
            bad = prev[:-1] + ("0" if prev[-1] != "0" else "1")
            gpu_ctx.set_template(0, prefix, bytes.fromhex(bad), n_int, 16)
            rc, res = gpu_ctx.search_min(0, n0, 1, max_nonce=n0 + 1)
            assert rc == 1


def test_rollover_order(gpu_ctx):
    """ts <- ts+1, nonce <- 0 when the nonce space is exhausted; lexicographic (ts, nonce) minimum."""
    for version in (1, 2):
        prefix = O.t0_prefix(3.0, version)
        rule = make_rule(3.0, O.T0_PREV_HASH)
        first = O.min_valid_nonce(prefix, rule)
        cap = max(first // 3, 256)
        want = O.first_valid_with_rollover(prefix, rule, 200, cap)
        assert want is not None and want[0] > O.T0_TIMESTAMP
        _set(gpu_ctx, 0, prefix, rule)
        rc, res = gpu_ctx.search_min(0, max_nonce=cap, max_rollovers=200)
        assert rc == 0 and (res.timestamp, res.nonce) == want
        assert res.rollovers == want[0] - O.T0_TIMESTAMP
        assert bytes(res.digest) == O.digest_of(O.set_timestamp(prefix, want[0]), want[1])
        rc, res = gpu_ctx.search_min(0, max_nonce=cap, max_rollovers=want[0] - O.T0_TIMESTAMP - 1)
        assert rc == 1 and res.found == 0


def test_two_queues_on_one_gpu_same_answer(sweep_fixture):
    """The multi-device tile scheduler, exercised with two device slots on GPU 0: same exact minimum,
    same sweep set, same rollover answer as one device."""
    import upow_b200
    with upow_b200.Context(device_ids=[0, 0]) as ctx2:
        assert ctx2.n_devices == 2
        prefix = bytes.fromhex(sweep_fixture["prefix"])
        _set(ctx2, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
        ns = sweep_fixture["nonces"]
        for tile in (1 << 16, 1 << 20, 1 << 24, 0):
            rc, res = ctx2.search_min(0, 0, 1, tile_nonces=tile)
            assert (rc, res.nonce) == (0, 164867), tile
            rc, res = ctx2.search_min(0, ns[40] + 1, 1, tile_nonces=tile)
            assert (rc, res.nonce) == (0, ns[41]), tile
        got, stats = ctx2.sweep(0, 0, 1 << 30)
        assert got == [n for n in ns if n < (1 << 30)]
        assert stats.per_device_hashes[0] > 0 and stats.per_device_hashes[1] > 0
        # strided + rollover through the scheduler
        p3 = O.t0_prefix(3.0)
        r3 = make_rule(3.0, O.T0_PREV_HASH)
        first = O.min_valid_nonce(p3, r3)
        cap = max(first // 3, 256)
        want = O.first_valid_with_rollover(p3, r3, 200, cap)
        _set(ctx2, 1, p3, r3)
        rc, res = ctx2.search_min(1, max_nonce=cap, max_rollovers=200, tile_nonces=1 << 16)
        assert rc == 0 and (res.timestamp, res.nonce) == want
        _set(ctx2, 2, O.t0_prefix(4.6), make_rule(4.6, O.T0_PREV_HASH))
        rc, res = ctx2.search_min(2, 3, 8, tile_nonces=1 << 16)
        want_n, _ = c_oracle.min_valid_nonce(O.t0_prefix(4.6), make_rule(4.6, O.T0_PREV_HASH), 3, 8)
        assert (rc, res.nonce) == (0, want_n)


def test_two_physical_gpus_same_answers(sweep_fixture):
    """The same checks as test_two_queues_on_one_gpu_same_answer on two REAL devices (finds travel through peer
    memory over NVLink); skipped on a one-GPU box."""
    import upow_b200
    try:
        ctx2 = upow_b200.Context(device_ids=[0, 1])
    except upow_b200.UpowError:
        pytest.skip("needs two GPUs")
    with ctx2:
        prefix = bytes.fromhex(sweep_fixture["prefix"])
        _set(ctx2, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
        ns = sweep_fixture["nonces"]
        for tile in (1 << 20, 0):
            rc, res = ctx2.search_min(0, 0, 1, tile_nonces=tile)
            assert (rc, res.nonce) == (0, 164867)
            assert res.per_device_hashes[0] > 0 and res.per_device_hashes[1] > 0
        got, stats = ctx2.sweep(0, 0, 1 << 30)
        assert got == [n for n in ns if n < (1 << 30)]
        _set(ctx2, 1, O.t0_prefix(6.3), make_rule(6.3, O.T0_PREV_HASH))
        rc, res = ctx2.search_min(1)
        assert (rc, res.nonce) == (0, 9436117)
        d = 4.0
        templates = [(O.batch_template(i, d), bytes.fromhex(O.T0_PREV_HASH), 4, 16) for i in range(64)]
        items, _ = ctx2.search_batch(templates)
        for i, (found, nonce, digest) in enumerate(items):
            assert found and nonce == c_oracle.min_valid_nonce(templates[i][0], make_rule(d, O.T0_PREV_HASH))[0]
        assert ctx2.device_info(0).peer_devices == 1


@pytest.mark.parametrize("n_slots", [3, 5])
def test_uneven_lgroup_classes_same_answers(gpu_ctx, sweep_fixture, n_slots):
    """2048 lgroups do not divide evenly over 3 or 5 device slots: the classes differ in size by one, every nonce
    still belongs to exactly one slot, and every answer equals the single-device / oracle answer."""
    import upow_b200
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    ns = sweep_fixture["nonces"]
    with upow_b200.Context(device_ids=[0] * n_slots) as ctx:
        _set(ctx, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
        for start_idx in (0, 17, 400):
            start = ns[start_idx - 1] + 1 if start_idx else 0
            rc, res = ctx.search_min(0, start)
            assert (rc, res.nonce) == (0, ns[start_idx])
            assert all(res.per_device_hashes[i] > 0 for i in range(n_slots))
        got, _ = ctx.sweep(0, 0, 1 << 27)
        assert got == [n for n in ns if n < (1 << 27)]
        p3, r3 = O.t0_prefix(3.0, 2), make_rule(3.0, O.T0_PREV_HASH)
        cap = max(O.min_valid_nonce(p3, r3) // 3, 256)
        want = O.first_valid_with_rollover(p3, r3, 200, cap)
        _set(ctx, 1, p3, r3)
        rc, res = ctx.search_min(1, max_nonce=cap, max_rollovers=200)
        assert rc == 0 and (res.timestamp, res.nonce) == want
        d = 4.0
        templates = [(O.batch_template(i, d), bytes.fromhex(O.T0_PREV_HASH), 4, 16) for i in range(40)]
        assert ctx.search_batch(templates)[0] == gpu_ctx.search_batch(templates)[0]


def test_batch_on_two_queues_matches_one(gpu_ctx):
    """The batch engine with two device slots on GPU 0: each takes half of every template's lgroups and the
    finds reach the other slot's limit array; answers equal the single-queue run and the oracle."""
    import upow_b200
    d = 4.0
    rule = make_rule(d, O.T0_PREV_HASH)
    templates = [(O.batch_template(i, d) if i % 3 else O.t0_prefix(d, 2, O.T0_TIMESTAMP + i), bytes.fromhex(O.T0_PREV_HASH), 4, 16)
                 for i in range(160)]
    one, _ = gpu_ctx.search_batch(templates)
    with upow_b200.Context(device_ids=[0, 0]) as ctx2:
        two, stats = ctx2.search_batch(templates)
        assert stats.per_device_hashes[0] > 0 and stats.per_device_hashes[1] > 0
        rc, idx, nonce, digest, _ = ctx2.search_any(templates)
    assert one == two
    for i in range(0, 160, 9):
        assert one[i][1] == c_oracle.min_valid_nonce(templates[i][0], rule)[0]
    rc1, idx1, nonce1, digest1, _ = gpu_ctx.search_any(templates)
    assert (rc, idx, nonce, digest) == (rc1, idx1, nonce1, digest1) and rc == 0


def test_without_peer_access_results_stay_exact(monkeypatch, sweep_fixture):
    """No peer access between the device slots (forced here with UPOW_NO_P2P=all): finds are not published GPU to
    GPU, the batch engine relays them through the host once per chunk of rounds, single searches simply run each
    slot's launch to its own end.  Slower, but every answer is still the exact minimum."""
    import upow_b200
    monkeypatch.setenv("UPOW_NO_P2P", "all")
    with upow_b200.Context(device_ids=[0, 0]) as ctx2:
        assert ctx2.device_info(0).peer_devices == 0
        prefix = bytes.fromhex(sweep_fixture["prefix"])
        _set(ctx2, 0, prefix, make_rule(5.0, O.T0_PREV_HASH))
        rc, res = ctx2.search_min(0)
        assert (rc, res.nonce) == (0, 164867)
        d = 4.0
        rule = make_rule(d, O.T0_PREV_HASH)
        templates = [(O.batch_template(i, d), bytes.fromhex(O.T0_PREV_HASH), 4, 16) for i in range(200)]
        items, _ = ctx2.search_batch(templates)
        for i in range(0, 200, 7):
            assert items[i][0] and items[i][1] == c_oracle.min_valid_nonce(templates[i][0], rule)[0]
        rc, idx, nonce, digest, _ = ctx2.search_any(templates)
    monkeypatch.delenv("UPOW_NO_P2P")
    with upow_b200.Context(1) as ctx1:
        assert ctx1.search_batch(templates)[0] == items
        assert ctx1.search_any(templates)[:4] == (rc, idx, nonce, digest)


@pytest.mark.parametrize("mix", ["v2", "mixed"])
def test_batch_4096_templates_v2_and_mixed(gpu_ctx, mix):
    """4096 templates of the 108-byte layout (and a mix of both layouts) at an oracle-checkable difficulty: the
    batch kernel's hand-specialised 108-byte loop returns every template's minimum."""
    d = 3.0
    rule = make_rule(d, O.T0_PREV_HASH)
    templates = []
    for i in range(4096):
        if mix == "v2" or i % 2:
            prefix = O.t0_prefix(d, 2, O.T0_TIMESTAMP + i)
        else:
            prefix = O.batch_template(i, d)
        templates.append((prefix, bytes.fromhex(O.T0_PREV_HASH), rule.int_difficulty, rule.charset_count))
    items, stats = gpu_ctx.search_batch(templates)
    for i, (found, nonce, digest) in enumerate(items):
        want, _ = c_oracle.min_valid_nonce(templates[i][0], rule, threads=1)
        assert found and nonce == want, (mix, i)
        assert digest == O.digest_of(templates[i][0], want)


def test_batch_templates_match_oracle(gpu_ctx):
    """config 5 shape at test size: per-template midstates, minimum nonce of each template."""
    d = 4.0
    templates, wants = [], []
    for i in range(96):
        version = 2 if i % 5 == 4 else 1
        if version == 1:
            prefix = O.batch_template(i, d)
        else:
            prefix = O.t0_prefix(d, 2, O.T0_TIMESTAMP + i)
        rule = make_rule(d if i % 7 else 4.4, O.T0_PREV_HASH)
        templates.append((prefix, bytes.fromhex(O.T0_PREV_HASH), rule.int_difficulty, rule.charset_count))
        wants.append(c_oracle.min_valid_nonce(prefix, rule)[0])
    items, stats = gpu_ctx.search_batch(templates)
    for i, ((found, nonce, digest), want) in enumerate(zip(items, wants)):
        assert found and nonce == want, i
        assert digest == O.digest_of(templates[i][0], want)
    assert stats.hashes_done > 0
    # bounded nonce space: templates whose minimum lies beyond max_nonce report not-found
    cap = sorted(wants)[len(wants) // 2]
    items, _ = gpu_ctx.search_batch(templates, max_nonce=cap)
    for (found, nonce, _), want in zip(items, wants):
        assert (found and nonce == want) if want < cap else not found
    assert gpu_ctx.search_batch([])[0] == []


def test_time_limit_and_cancel(gpu_ctx):
    """An unreachable difficulty: the time limit returns 1 (the reference's search_complete=False), cancel returns 2."""
    prefix = O.t0_prefix(16.0)
    _set(gpu_ctx, 0, prefix, make_rule(16.0, O.T0_PREV_HASH))
    t0 = time.time()
    rc, res = gpu_ctx.search_min(0, max_rollovers=100000, time_limit_s=0.5)
    assert rc == 1 and res.found == 0 and 0.4 < time.time() - t0 < 5
    assert res.hashes_done > 0
    timer = threading.Timer(0.3, gpu_ctx.cancel)
    timer.start()
    t0 = time.time()
    rc, res = gpu_ctx.search_min(0, max_rollovers=100000)
    assert rc == 2 and time.time() - t0 < 5
    # the context is reusable afterwards
    _set(gpu_ctx, 0, O.t0_prefix(4.0), make_rule(4.0, O.T0_PREV_HASH))
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 62804)


def test_cancel_before_search_start_is_honoured(gpu_ctx):
    """The cancel flag is sticky: a cancel that lands between "template fetched" and "search started" (the
    stale-work watcher runs in another thread) is not wiped by the next call -- that call returns CANCELLED at
    once, and consumes the flag.  The reference's equivalent: the parent terminates its workers, miner.py:149-156."""
    _set(gpu_ctx, 0, O.t0_prefix(16.0), make_rule(16.0, O.T0_PREV_HASH))
    gpu_ctx.cancel()
    t0 = time.time()
    rc, res = gpu_ctx.search_min(0, max_rollovers=100000)
    assert rc == 2 and res.hashes_done == 0 and time.time() - t0 < 0.5
    # consumed: the next call runs
    _set(gpu_ctx, 0, O.t0_prefix(4.0), make_rule(4.0, O.T0_PREV_HASH))
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 62804)
    # a pending cancel can be dropped explicitly (what miner.main does after fetching a fresh template)
    gpu_ctx.cancel()
    gpu_ctx.clear_cancel()
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 62804)
    # sweeps and batches honour a pending cancel too
    gpu_ctx.cancel()
    import upow_b200
    buf_rc = gpu_ctx._L.upow_sweep(gpu_ctx._h, 0, 0, 1 << 20, None, 0, __import__("ctypes").byref(__import__("ctypes").c_uint64()), None)
    assert buf_rc == upow_b200.UPOW_CANCELLED
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 62804)


def test_cancel_stops_a_running_batch_and_a_running_sweep(gpu_ctx):
    """upow_cancel from another thread while the batch engine (rounds queued on the device) or a full sweep is
    running: the call returns CANCELLED long before its work would have ended, and the context stays usable."""
    import upow_b200
    r16 = make_rule(16.0, O.T0_PREV_HASH)
    temps = [(O.t0_prefix(16.0, 1, O.T0_TIMESTAMP + i), bytes.fromhex(O.T0_PREV_HASH), 16, 16) for i in range(256)]
    timer = threading.Timer(0.15, gpu_ctx.cancel)
    t0 = time.time()
    timer.start()
    with pytest.raises(upow_b200.UpowError) as e:        # 256 x 2^32 hashes would take a minute
        gpu_ctx.search_batch(temps)
    timer.join()
    assert e.value.code == upow_b200.UPOW_CANCELLED and time.time() - t0 < 3
    _set(gpu_ctx, 0, O.t0_prefix(16.0), r16)
    timer = threading.Timer(0.05, gpu_ctx.cancel)
    timer.start()
    with pytest.raises(upow_b200.UpowError) as e:
        gpu_ctx.sweep(0, 0, 1 << 32)
    timer.join()
    assert e.value.code == upow_b200.UPOW_CANCELLED
    _set(gpu_ctx, 0, O.t0_prefix(4.0), make_rule(4.0, O.T0_PREV_HASH))
    rc, res = gpu_ctx.search_min(0)
    assert (rc, res.nonce) == (0, 62804)
    items, _ = gpu_ctx.search_batch([(O.batch_template(i, 4.0), bytes.fromhex(O.T0_PREV_HASH), 4, 16) for i in range(8)])
    assert all(f for f, _, _ in items)


def test_backwards_rollover_order(gpu_ctx):
    """rollover_backwards: timestamps ts, ts-1, ts-2, ... (what a live miner needs: the node rejects timestamps
    in the future, upow/manager.py:456-460).  First valid (rollover index, nonce) in that order."""
    for version in (1, 2):
        prefix = O.t0_prefix(3.0, version)
        rule = make_rule(3.0, O.T0_PREV_HASH)
        cap = max(O.min_valid_nonce(prefix, rule) // 3, 256)
        want = None
        for k in range(200):
            n = O.min_valid_nonce(O.set_timestamp(prefix, O.T0_TIMESTAMP - k), rule, 0, 1, cap)
            if n is not None:
                want = (k, O.T0_TIMESTAMP - k, n)
                break
        assert want is not None and want[0] > 0
        _set(gpu_ctx, 0, prefix, rule)
        rc, res = gpu_ctx.search_min(0, max_nonce=cap, max_rollovers=200, rollover_backwards=True)
        assert rc == 0 and (res.rollovers, res.timestamp, res.nonce) == want
        assert bytes(res.digest) == O.digest_of(O.set_timestamp(prefix, want[1]), want[2])
        rc, res = gpu_ctx.search_min(0, max_nonce=cap, max_rollovers=want[0] - 1, rollover_backwards=True)
        assert rc == 1


def test_progress_counter(gpu_ctx):
    _set(gpu_ctx, 0, O.t0_prefix(16.0), make_rule(16.0, O.T0_PREV_HASH))
    rc, res = gpu_ctx.search_min(0, max_nonce=1 << 28, max_rollovers=3)
    assert rc == 1 and gpu_ctx.progress() == res.hashes_done == 4 << 28


def test_hashes_done_tracks_the_answer(gpu_ctx):
    """Early exit granularity: a search stops within a few H values of the find on every warp, so the hashes
    evaluated stay close to nonce + 1 (BASELINE configs[1])."""
    for d, want in ((6.3, 9436117), (5.0, 164867)):
        _set(gpu_ctx, 0, O.t0_prefix(d), make_rule(d, O.T0_PREV_HASH))
        rc, res = gpu_ctx.search_min(0)
        assert (rc, res.nonce) == (0, want)
        assert res.hashes_done >= want + 1 - 65536 * 32     # lanes of the last H values may not all have run
        assert res.hashes_done <= 1.35 * (want + 1) + (1 << 22), (d, res.hashes_done)


def test_argument_errors(gpu_ctx):
    import upow_b200
    with pytest.raises(upow_b200.UpowError) as e:
        gpu_ctx.set_template(0, b"x" * 100, bytes(32), 4, 16)
    assert e.value.code == -1
    with pytest.raises(upow_b200.UpowError):
        gpu_ctx.set_template(99, O.t0_prefix(4.0), bytes(32), 4, 16)
    with pytest.raises(upow_b200.UpowError):
        gpu_ctx.search_min(63)          # slot never set
    with pytest.raises(upow_b200.UpowError):
        gpu_ctx.search_min(0, 0, 0)     # stride 0
    good = (O.t0_prefix(4.0), bytes(32), 4, 16)
    for bad in ((b"x" * 90, bytes(32), 4, 16), (O.t0_prefix(4.0), bytes(32), 4, 17), (O.t0_prefix(4.0), bytes(32), 64, 8)):
        with pytest.raises(upow_b200.UpowError) as e:
            gpu_ctx.search_batch([good] * 3 + [bad])
        assert e.value.code == -1 and "template 3" in str(e.value)
        with pytest.raises(upow_b200.UpowError):
            gpu_ctx.search_any([bad])


def test_linearity_of_hash_counts(gpu_ctx):
    """Size-independent property: hashes_done of disjoint sweeps adds up to the sweep of the union."""
    _set(gpu_ctx, 0, O.t0_prefix(5.0), make_rule(5.0, O.T0_PREV_HASH))
    a, sa = gpu_ctx.sweep(0, 0, 1 << 26)
    b, sb = gpu_ctx.sweep(0, 1 << 26, 1 << 27)
    c, sc = gpu_ctx.sweep(0, 0, 1 << 27)
    assert a + b == c and sa.hashes_done + sb.hashes_done == sc.hashes_done == 1 << 27


def test_miner_mirror_end_to_end_on_gpu(monkeypatch, capsys, mine_block_vectors):
    import sys, types
    from upow_b200 import miner
    posted = {}

    class Resp:
        def json(self):
            return {"ok": True}

    fake = types.ModuleType("requests")
    fake.post = lambda url, json=None, timeout=None: (posted.update(json), Resp())[1]
    monkeypatch.setitem(sys.modules, "requests", fake)
    miner.set_context(None)
    for v in mine_block_vectors[:6]:
        posted.clear()
        monkeypatch.setattr(sys, "argv", ["miner.py", v["wallet"], "1"])
        monkeypatch.setattr(miner, "timestamp", lambda ts=v["timestamp"]: ts)
        params = {k: (dict(x) if isinstance(x, dict) else x) for k, x in v["mining_params"].items()}
        with pytest.raises(SystemExit):
            miner.mine_block(0, 1, params)
        assert posted["block_content"] == v["block_content"]
        assert posted["block_no"] == v["posted_block_no"]
    capsys.readouterr()
