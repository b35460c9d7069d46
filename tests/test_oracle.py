"""The oracle against everything that pins it: FIPS KATs, hashlib, the golden vectors produced
by the reference's own mine_block / check_block_is_valid, and (when /root/reference is present)
the reference itself run verbatim."""
import hashlib
import os
import random

import pytest

from oracle import c_oracle, ref_loader
from oracle import upow_oracle as O

KATS = {
    b"": "e3b0c44298fc1c149afbf4c8996fb92427ae41e4649b934ca495991b7852b855",
    b"abc": "ba7816bf8f01cfea414140de5dae2223b00361a396177a9cb410ff61f20015ad",
    b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq":
        "248d6a61d20638b8e5c026930c3e6039a33ce45964ff2167f6ecedd419db06c1",
}


def test_fips_kats_c_oracle_both_backends():
    for shani in (True, False):
        c_oracle.set_shani(shani)
        for msg, want in KATS.items():
            assert c_oracle.sha256(msg).hex() == want
    c_oracle.set_shani(True)


def test_python_compress_matches_hashlib():
    rng = random.Random(1)
    for version in (1, 2):
        for _ in range(5):
            prefix = rng.randbytes(134 if version == 1 else 104)
            nonce = rng.getrandbits(32)
            assert O.tail_digest(prefix, nonce) == O.digest_of(prefix, nonce)
            assert c_oracle.tail_digest(prefix, nonce) == O.digest_of(prefix, nonce)


def test_c_oracle_random_messages():
    rng = random.Random(2)
    for n in (0, 1, 55, 56, 63, 64, 65, 108, 119, 120, 138, 1000):
        m = rng.randbytes(n)
        assert c_oracle.sha256(m) == hashlib.sha256(m).digest()


def test_t0_template_constants():
    # SURVEY 8(d) / Appendix A.1
    assert O.T0_PREV_HASH.startswith("f6652fdb") and O.T0_PREV_HASH.endswith("053745dc")
    assert O.calculate_merkle_root(O.T0_TX_HASHES).startswith("6afcaf4a")
    assert O.T0_ADDRESS_V2 == "DZbXJHETzx7yVQoAZBzt2zdq26c2aP6nfoS6jsR9ZCd3T"
    assert O.calculate_merkle_root([]) == KATS[b""]
    assert len(O.t0_prefix(4.0, 1)) == 134 and len(O.t0_prefix(4.0, 2)) == 104


def test_oracle_reproduces_reference_mine_block(mine_block_vectors):
    """Every golden vector (reference mine_block run verbatim) from the restated loop."""
    for v in mine_block_vectors:
        p = v["mining_params"]
        lb = p["last_block"]
        prev = lb.get("hash", O.GENESIS_PREV_HASH)
        rule = O.Rule(p["difficulty"], prev)
        wallet = O.string_to_bytes(v["wallet"])
        prefix = O.build_prefix(prev, wallet, O.calculate_merkle_root(p["pending_transactions_hashes"]),
                                v["timestamp"], p["difficulty"])
        assert prefix.hex() == v["block_content"][:-8], v["name"]
        if v["nonce"] < 3_000_000:
            n = O.min_valid_nonce(prefix, rule, v["start_nonce"], v["nonce_increment"])
        else:
            n, _ = c_oracle.min_valid_nonce(prefix, rule, v["start_nonce"], v["nonce_increment"])
        assert n == v["nonce"], v["name"]
        assert O.digest_of(prefix, n).hex() == v["digest"]
        cn, _ = c_oracle.min_valid_nonce(prefix, rule, v["start_nonce"], v["nonce_increment"])
        assert cn == v["nonce"], v["name"]
        assert v["node_accepts"] is True
        assert v["posted_block_no"] == lb.get("id", 0) + 1


def test_survey_appendix_a_values(mine_block_vectors):
    by = {v["name"]: v for v in mine_block_vectors}
    assert by["T0-v1 d=4.0"]["nonce"] == 62804
    assert by["T0-v1 d=4"]["nonce"] == 62804
    assert by["T0-v1 d=5.3"]["nonce"] == 2373524
    assert by["T0-v1 d=6.3"]["nonce"] == 9436117
    assert by["T0-v1 d=6.3"]["digest"] == "3745dc106adcf74ca0691a8f6031a651ef63add4c44d78962e83dfc36c884343"
    assert by["T0-v2 d=4.0"]["nonce"] == 86018
    assert by["T0-v2 d=5.3"]["nonce"] == 39344


def test_node_verifier_restatement(verifier_vectors):
    for v in verifier_vectors:
        assert O.node_check_block_is_valid(v["block_content"], v["difficulty"], v["last_block"]) == v["valid"]
        # miner-side rule (float) agrees with the node-side rule (Decimal) on the same block
        rule = O.Rule(v["difficulty"], v["last_block"]["hash"])
        assert rule.is_valid_block(bytes.fromhex(v["block_content"])) == v["valid"]


def test_rule_counts_float_vs_decimal():
    """charset counts for .0-.9: 16,15,13,12,10,8,7,5,4,2."""
    want = [16, 15, 13, 12, 10, 8, 7, 5, 4, 2]
    for n in range(0, 20):
        for k in range(10):
            d = float(f"{n}.{k}")
            r = O.Rule(d, O.T0_PREV_HASH)
            assert r.charset_count == want[k], d
            assert r.int_difficulty == n
            assert int(d * 10) == n * 10 + k


def test_rule_quirks():
    r = O.Rule(0.5, O.T0_PREV_HASH)  # N == 0: s[-0:] is the whole string
    assert r.target_prefix == O.T0_PREV_HASH
    r = O.Rule(6, O.T0_PREV_HASH)    # int difficulty
    assert r.target_prefix == "3745dc" and r.charset_count == 16


def test_sweep_fixture_consistency(sweep_fixture):
    ns = sweep_fixture["nonces"]
    assert len(ns) == sweep_fixture["count"] == 4209
    assert ns == sorted(ns)
    assert O.nonce_set_hash(ns) == sweep_fixture["sha256_sorted_u32le"] \
        == "5ca83877d72644292a359a6a2864d9411523a9c01a2f2a0ec6aea1ee3c9d742f"
    assert ns[:8] == [164867, 2752366, 3149421, 5107810, 5554151, 6120560, 6658588, 7305164]
    assert ns[-1] == 4293591526
    assert sweep_fixture["cumulative_per_2^28"] == [268, 525, 783, 1072, 1338, 1575, 1829, 2122, 2359, 2632, 2892,
                                                    3135, 3412, 3694, 3943, 4209]
    # spot-check with hashlib: a sample of members is valid, their neighbours are not members unless listed
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    rule = O.Rule(5.0, O.T0_PREV_HASH)
    members = set(ns)
    for n in ns[::97]:
        assert rule.is_valid_block(prefix + n.to_bytes(4, "little"))
        for m in (n - 1, n + 1):
            if 0 <= m < 2 ** 32:
                assert rule.is_valid_block(prefix + m.to_bytes(4, "little")) == (m in members)


def test_sweep_slice_oracles_agree(sweep_fixture):
    """hashlib loop == C oracle == fixture on a slice."""
    prefix = bytes.fromhex(sweep_fixture["prefix"])
    rule = O.Rule(5.0, O.T0_PREV_HASH)
    lo, hi = 5_000_000, 5_000_000 + (1 << 20)
    want = [n for n in sweep_fixture["nonces"] if lo <= n < hi]
    assert O.all_valid_nonces(prefix, rule, lo, hi) == want
    got, hashes = c_oracle.all_valid_nonces(prefix, rule, lo, hi)
    assert got == want and hashes == hi - lo


def test_rollover_oracle():
    prefix = O.t0_prefix(3.0)
    rule = O.Rule(3.0, O.T0_PREV_HASH)
    # restrict the nonce space so the first timestamps have no solution
    first = O.min_valid_nonce(prefix, rule)
    res = O.first_valid_with_rollover(prefix, rule, max_rollovers=50, max_nonce=max(first // 4, 1))
    assert res is not None
    ts, n = res
    assert rule.is_valid_block(O.set_timestamp(prefix, ts) + n.to_bytes(4, "little"))
    for t in range(O.T0_TIMESTAMP, ts):
        assert O.min_valid_nonce(O.set_timestamp(prefix, t), rule, 0, 1, max(first // 4, 1)) is None


@pytest.mark.skipif(not ref_loader.reference_available(), reason="reference tree only exists in the build container")
def test_reference_verbatim_matches_oracle():
    rng = random.Random(7)
    for d in (2.0, 3.3, 4.0):
        prev = rng.randbytes(32).hex()
        wallet = rng.randbytes(64).hex()
        txs = sorted(rng.randbytes(32).hex() for _ in range(2))
        ts = 1_750_000_000 + rng.randrange(1000)
        params = {"difficulty": d, "last_block": {"hash": prev, "id": 7}, "pending_transactions_hashes": txs}
        out = ref_loader.run_mine_block(wallet, params, ts)
        prefix = O.build_prefix(prev, O.string_to_bytes(wallet), O.calculate_merkle_root(txs), ts, d)
        n = O.min_valid_nonce(prefix, O.Rule(d, prev))
        assert out["block_content"] == (prefix + n.to_bytes(4, "little")).hex()
        assert ref_loader.check_block_is_valid(out["block_content"], d, {"hash": prev, "id": 7})
        # the loop-shaped port that bench.py times as the CPU arm returns what the reference returns, also as worker (k, W)
        assert O.miner_loop(prefix, O.Rule(d, prev)) == n
        out3 = ref_loader.run_mine_block(wallet, params, ts, 2, 3)
        n3 = int.from_bytes(bytes.fromhex(out3["block_content"])[-4:], "little")
        assert O.miner_loop(prefix, O.Rule(d, prev), 2, 3) == n3 == O.min_valid_nonce(prefix, O.Rule(d, prev), 2, 3)


def test_miner_loop_port_matches_golden_vectors(mine_block_vectors):
    """miner_loop (the reference's loop shape, timed by bench.py as the CPU arm) against blocks the reference mined."""
    calls = []
    for v in mine_block_vectors[:12]:
        p = v["mining_params"]
        prev = p["last_block"].get("hash", O.GENESIS_PREV_HASH)
        prefix = bytes.fromhex(v["block_content"])[:-4]
        assert O.miner_loop(prefix, O.Rule(p["difficulty"], prev), v["start_nonce"], v["nonce_increment"],
                            progress=lambda n, e: calls.append(n)) == v["nonce"]
    assert O.miner_loop(O.t0_prefix(16.0), O.Rule(16.0, O.T0_PREV_HASH), 0, 1, 1000) is None


def test_merkle_tree_restatement_against_reference_vectors():
    import json, os
    from conftest import GOLDEN
    cases = json.load(open(os.path.join(GOLDEN, "merkle_tree_vectors.json")))
    assert len(cases) == 5
    for c in cases:
        assert O.transactions_merkle_tree(c["txs"]) == c["merkle_tree"]
