"""BASELINE configs 4 and 5 on the GPU.

Each config is covered three ways: at an oracle-checkable size with identical semantics (bit-exact vs the C
oracle run inside the test); at FULL size against fixtures the C oracle wrote once in the build container
(oracle/gen_golden.py --cfg4 --cfg5: every nonce below the config-4 answer at 9.0 walked and rejected, the
first valid nonce of all 4096 config-5 templates; log in tests/golden/cfg_pins.log); and, where no CPU could
walk the space (9.3: 1e11 hashes), through properties -- the answer re-hashed with hashlib and checked against
the reference rule, and the same answer from independent device implementations."""
import hashlib
import json
import os
import random
import struct

import pytest

from oracle import c_oracle
from oracle import upow_oracle as O

pytestmark = pytest.mark.gpu
PREV = bytes.fromhex(O.T0_PREV_HASH)


def _oracle_rollover(prefix, rule, max_nonce, max_rollovers):
    ts0 = O.get_timestamp(prefix)
    for r in range(max_rollovers + 1):
        n, _ = c_oracle.min_valid_nonce(O.set_timestamp(prefix, ts0 + r), rule, 0, 1, max_nonce)
        if n is not None:
            return ts0 + r, n
    return None


@pytest.mark.parametrize("version,difficulty", [(1, 7.0), (2, 6.6)])
def test_config4_semantics_at_oracle_size(version, difficulty):
    """Rollover search (ts <- ts+1, nonce <- 0) over a nonce space cut to 2^22 so that dozens of rollovers
    happen: lexicographic (ts, nonce) minimum, identical for 1 queue, 2 queues and small tiles."""
    import upow_b200
    prefix = O.t0_prefix(difficulty, version)
    rule = O.Rule(difficulty, O.T0_PREV_HASH)
    want = _oracle_rollover(prefix, rule, 1 << 22, 400)
    assert want is not None and want[0] > O.T0_TIMESTAMP + 3
    for ids, tile in (([0], 0), ([0, 0], 1 << 20), ([0, 0], 0)):
        with upow_b200.Context(device_ids=ids) as ctx:
            ctx.set_template(0, prefix, PREV, rule.int_difficulty, rule.charset_count)
            rc, res = ctx.search_min(0, max_nonce=1 << 22, max_rollovers=400, tile_nonces=tile)
            assert rc == 0 and (res.timestamp, res.nonce) == want, (ids, tile)
            assert bytes(res.digest) == O.digest_of(O.set_timestamp(prefix, want[0]), want[1])


def test_config4_full_size_against_the_cpu_oracle_fixture():
    """T0-v1 at difficulty 9.0: the first valid nonce in the order of miner.py:86-89, as the C oracle found it
    by walking all 2.9e9 nonces below it -- on one queue, two queues and with small launch granules."""
    import upow_b200
    from conftest import GOLDEN
    fx = json.load(open(os.path.join(GOLDEN, "cfg4_t0v1_d9.json")))
    d = fx["difficulty"]
    prefix = O.t0_prefix(d)
    rule = O.Rule(d, O.T0_PREV_HASH)
    for ids, tile in (([0], 0), ([0, 0], 0), ([0, 0], 1 << 28)):
        with upow_b200.Context(device_ids=ids) as ctx:
            ctx.set_template(0, prefix, PREV, rule.int_difficulty, rule.charset_count)
            rc, res = ctx.search_min(0, max_rollovers=16, tile_nonces=tile)
            assert rc == 0 and (res.timestamp, res.nonce, res.digest_hex) == (fx["timestamp"], fx["nonce"], fx["digest"]), (ids, tile)
            assert res.rollovers == 0 and res.hashes_done >= fx["nonce"] - (1 << 21)


def test_config4_full_size_consistency():
    """Difficulty 9.3 (charset-restricted 10th nibble) with rollover: ~1e11 hashes.  Same (ts, nonce, digest)
    from the plain-CUDA kernel, the tuned kernel and the two-queue scheduler; valid under the reference rule."""
    import upow_b200
    d = 9.3
    prefix = O.t0_prefix(d)
    rule = O.Rule(d, O.T0_PREV_HASH)
    answers = []
    for ids, variant, tile in (([0], None, 0), ([0, 0], None, 1 << 28), ([0], 0, 0)):
        with upow_b200.Context(device_ids=ids) as ctx:
            if variant is not None:
                ctx.variant = variant
            ctx.set_template(0, prefix, PREV, rule.int_difficulty, rule.charset_count)
            rc, res = ctx.search_min(0, max_rollovers=4096, tile_nonces=tile)
            assert rc == 0
            answers.append((res.timestamp, res.nonce, res.digest_hex))
    assert answers[0] == answers[1] == answers[2]
    ts, nonce, digest = answers[0]
    block = O.set_timestamp(prefix, ts) + nonce.to_bytes(4, "little")
    assert hashlib.sha256(block).hexdigest() == digest
    assert rule.is_valid_block(block)
    assert O.node_check_block_is_valid(block.hex(), d, {"hash": O.T0_PREV_HASH})
    assert digest.startswith(O.T0_PREV_HASH[-9:]) and digest[9] in "0123456789ab"


def test_config5_4096_templates():
    """4096 templates at difficulty 6.0 (6.9e10 expected hashes): EVERY minimum equals the C oracle's
    (tests/golden/cfg5_min_nonces.u32, one sequential walk per template), all digests are hashlib's; a subset
    equals what the single-template search kernel finds."""
    import upow_b200
    from conftest import GOLDEN
    d = 6.0
    rule = O.Rule(d, O.T0_PREV_HASH)
    prefixes = [O.batch_template(i, d) for i in range(4096)]
    assert len(set(prefixes)) == 4096
    raw = open(os.path.join(GOLDEN, "cfg5_min_nonces.u32"), "rb").read()
    want = list(struct.unpack("<4096I", raw))
    meta = json.load(open(os.path.join(GOLDEN, "cfg5_min_nonces.json")))
    assert hashlib.sha256(raw).hexdigest() == meta["sha256_u32le_in_template_order"]
    with upow_b200.Context(1) as ctx:
        items, stats = ctx.search_batch([(p, PREV, rule.int_difficulty, rule.charset_count) for p in prefixes])
        assert [n for _, n, _ in items] == want
        assert stats.hashes_done >= sum(want) and stats.hashes_done < 1.2 * (sum(want) + 4096)
        for (found, nonce, digest), p in zip(items, prefixes):
            assert found
            block = p + nonce.to_bytes(4, "little")
            assert hashlib.sha256(block).digest() == digest and rule.is_valid_block(block)
        rng = random.Random(5)
        for i in rng.sample(range(4096), 24):
            ctx.set_template(0, prefixes[i], PREV, rule.int_difficulty, rule.charset_count)
            rc, res = ctx.search_min(0)
            assert rc == 0 and res.nonce == items[i][1] and bytes(res.digest) == items[i][2], i
