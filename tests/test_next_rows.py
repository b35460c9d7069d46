"""SURVEY 8(f) "next" rows built so far: the template multiplier (row 1) and the stale-work abort (row 2).
Host logic runs on the CPU; the GPU legs race templates through upow_search_any and check the winner against
the oracle and against the node-side rules restated from the reference."""
import hashlib
import threading
import time

import pytest

from oracle import c_oracle, ref_loader
from oracle import upow_oracle as O
from upow_b200 import multiplier, stale


def _mining_info(difficulty, n_txs=3, last_ts=O.T0_TIMESTAMP - 20):
    txs = sorted(hashlib.sha256(b"tx%d" % i).hexdigest() for i in range(n_txs))
    return {"difficulty": difficulty,
            "last_block": {"hash": O.T0_PREV_HASH, "id": 1000, "timestamp": last_ts},
            "pending_transactions_hashes": txs}


def node_merkle(tx_hashes):
    """upow/manager.py:365-378 for a list of tx *hashes* the way push_block resolves them: the node looks the
    transactions up, sorts their bytes and hashes each -- for an order-preserving subset of its own sorted
    list that equals hashing the concatenated tx hashes in the given order."""
    return hashlib.sha256(b"".join(bytes.fromhex(h) for h in tx_hashes)).hexdigest()


def test_tx_subsets_are_order_preserving_and_unique():
    txs = ["a", "b", "c", "d"]
    subs = list(multiplier.tx_subsets(txs, 100))
    assert subs[0] == ("a", "b", "c", "d") and len(subs) == 16 == len(set(subs))
    for s in subs:
        assert list(s) == sorted(s)
    assert [len(s) for s in subs[:5]] == [4, 3, 3, 3, 3]       # fullest first
    assert len(list(multiplier.tx_subsets(txs, 5))) == 5


def test_candidates_respect_the_node_rules():
    info = _mining_info(6.0, n_txs=3)
    now = O.T0_TIMESTAMP
    cands = multiplier.candidates(info, O.T0_ADDRESS_V1, now, 100)
    assert len(cands) == 100 and len({c.prefix for c in cands}) == 100
    for c in cands:
        # timestamp window of upow/manager.py:445-460
        assert info["last_block"]["timestamp"] < c.timestamp <= now
        # merkle rule of upow/manager.py:365-378 / miner.py:15-18
        assert c.merkle_root == node_merkle(c.tx_hashes)
        assert set(c.tx_hashes) <= set(info["pending_transactions_hashes"])
        # same bytes the reference builds for that (txs, timestamp)
        assert c.prefix == O.build_prefix(O.T0_PREV_HASH, O.string_to_bytes(O.T0_ADDRESS_V1), c.merkle_root,
                                          c.timestamp, 6.0)
    assert cands[0].timestamp == now and len(cands[0].tx_hashes) == 3
    assert multiplier.candidates(_mining_info(6.0, last_ts=now), O.T0_ADDRESS_V1, now, 10) == []   # empty window
    v2 = multiplier.candidates(info, O.T0_ADDRESS_V2, now, 4)
    assert all(len(c.prefix) == 104 for c in v2)


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_subset_merkle_equals_reference_merkle_tree():
    mgr = ref_loader.load_manager()
    txs_hex = sorted(hashlib.sha256(b"rawtx%d" % i).digest().hex() * 3 for i in range(5))   # fake tx hex blobs
    hashes = [hashlib.sha256(bytes.fromhex(t)).hexdigest() for t in txs_hex]               # main.py:692
    for subset in multiplier.tx_subsets(list(range(5)), 12):
        chosen_hex = [txs_hex[i] for i in subset]
        chosen_hashes = [hashes[i] for i in subset]
        assert mgr.get_transactions_merkle_tree(chosen_hex) == multiplier.calculate_merkle_root(chosen_hashes)


class _FakeCtx:
    def __init__(self):
        self.cancelled = 0

    def cancel(self):
        self.cancelled += 1


def test_stale_watcher_cancels_on_new_tip_only():
    ctx = _FakeCtx()
    tips = iter([1000, 1000, None, 1001, 1001])
    with stale.StaleWatcher(ctx, 1000, lambda: next(tips), period_s=0.02) as w:
        time.sleep(0.4)
    assert w.fired and ctx.cancelled == 1
    ctx2 = _FakeCtx()

    def flaky():
        raise OSError("node unreachable")
    with stale.StaleWatcher(ctx2, 1000, flaky, period_s=0.02) as w2:
        time.sleep(0.1)
    assert not w2.fired and ctx2.cancelled == 0


@pytest.mark.gpu
def test_search_any_matches_oracle_and_is_gpu_count_independent():
    import upow_b200
    d = 5.0
    rule = O.Rule(d, O.T0_PREV_HASH)
    prev = bytes.fromhex(O.T0_PREV_HASH)
    prefixes = [O.batch_template(i, d) for i in range(300)] + [O.t0_prefix(d, 2, O.T0_TIMESTAMP + i) for i in range(20)]
    temps = [(p, prev, rule.int_difficulty, rule.charset_count) for p in prefixes]
    mins = [c_oracle.min_valid_nonce(p, rule)[0] for p in prefixes]
    answers = []
    for ids in ([0], [0, 0], [0, 0, 0]):
        with upow_b200.Context(device_ids=ids) as ctx:
            rc, idx, nonce, digest, stats = ctx.search_any(temps)
            assert rc == 0
            answers.append((idx, nonce, digest))
            # the winner's nonce is that template's true minimum, and no template with a lower index has a
            # valid nonce at or below the winner's round boundary
            assert nonce == mins[idx]
            assert digest == O.digest_of(prefixes[idx], nonce)
    assert answers[0] == answers[1] == answers[2]
    idx, nonce, _ = answers[0]
    # a lower-index template with a find in the same (or an earlier) round would have won instead
    assert all(mins[i] is None or mins[i] > nonce for i in range(idx))
    with upow_b200.Context(1) as ctx:      # nothing to find: unreachable difficulty in a small nonce space
        r16 = O.Rule(16.0, O.T0_PREV_HASH)
        rc, *_ = ctx.search_any([(p, prev, r16.int_difficulty, r16.charset_count) for p in prefixes[:8]], 1 << 20)
        assert rc == 1
        assert ctx.search_any([])[0] == 1


@pytest.mark.gpu
def test_mine_any_produces_a_block_the_node_would_accept():
    import upow_b200
    info = _mining_info(6.5, n_txs=4)
    now = O.T0_TIMESTAMP
    with upow_b200.Context(1) as ctx:
        out = multiplier.mine_any(ctx, info, O.T0_ADDRESS_V1, now, max_templates=512)
    assert out is not None and out["templates"] == 320      # 20 timestamps x 2^4 transaction subsets
    block = bytes.fromhex(out["block_content"])
    assert hashlib.sha256(block).hexdigest() == out["digest"]
    assert O.node_check_block_is_valid(out["block_content"], 6.5, info["last_block"])        # manager.py:130-151
    ts = int.from_bytes(block[-10:-6], "little")
    assert info["last_block"]["timestamp"] < ts <= now and ts == out["timestamp"]           # manager.py:445-460
    assert block[96:128].hex() == node_merkle(out["txs"])                                   # manager.py:639-645
    assert out["block_no"] == 1001
    # the nonce is the minimum for the winning template
    cands = multiplier.candidates(info, O.T0_ADDRESS_V1, now, 512)
    want, _ = c_oracle.min_valid_nonce(cands[out["template_index"]].prefix, O.Rule(6.5, O.T0_PREV_HASH))
    assert out["nonce"] == want


@pytest.mark.gpu
def test_stale_watcher_aborts_a_running_search():
    import upow_b200
    rule = O.Rule(16.0, O.T0_PREV_HASH)
    with upow_b200.Context(1) as ctx:
        ctx.set_template(0, O.t0_prefix(16.0), bytes.fromhex(O.T0_PREV_HASH), rule.int_difficulty, rule.charset_count)
        t_new_block = time.time() + 0.3
        with stale.StaleWatcher(ctx, 1000, lambda: 1001 if time.time() > t_new_block else 1000, period_s=0.05) as w:
            t0 = time.time()
            rc, res = ctx.search_min(0, max_rollovers=100000)
            dt = time.time() - t0
        assert rc == upow_b200.UPOW_CANCELLED and w.fired and 0.25 < dt < 3.0
