"""Generate tests/golden/* from the reference itself -- TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):

    python -m oracle.gen_golden            # mine_block / verifier vectors (about a minute)
    python -m oracle.gen_golden --sweep    # + the 2^32 sweep fixture (minutes, C oracle)

* ``mine_block_vectors.json``: every entry is ``miner.mine_block(start, inc, params)``
  (miner.py:30-114) run VERBATIM with only ``requests.post`` and ``timestamp()``
  replaced, plus the verdict of ``upow.manager.check_block_is_valid``
  (upow/manager.py:130-151) on the resulting block, also run verbatim.
* ``verifier_vectors.json``: check_block_is_valid verdicts on near-miss blocks.
* ``merkle_tree_vectors.json``: get_transactions_merkle_tree (upow/manager.py:365-378) on random blobs.
* ``sweep_t0v1_d5.u32`` (+ ``.json``): all valid nonces of the full 2^32 sweep of
  template T0-v1 at difficulty 5.0, enumerated by the C
  oracle (oracle/upow_oracle.c, full-message SHA-256 per nonce); the script refuses
  to write it unless its fingerprint equals the one SURVEY A.3 recorded from a
  plain hashlib loop.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import random
import struct

from . import c_oracle, ref_loader
from . import upow_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
SURVEY_A3_HASH = "5ca83877d72644292a359a6a2864d9411523a9c01a2f2a0ec6aea1ee3c9d742f"
SURVEY_A3_COUNT = 4209


def _vector(name, wallet, params, ts, start=0, inc=1):
    out = ref_loader.run_mine_block(wallet, params, ts, start, inc)
    content = out["block_content"]
    raw = bytes.fromhex(content)
    last_block = dict(params["last_block"])
    verdict = ref_loader.check_block_is_valid(content, params["difficulty"], last_block) \
        if "hash" in last_block else True
    return {
        "name": name,
        "wallet": wallet,
        "mining_params": params,
        "timestamp": ts,
        "start_nonce": start,
        "nonce_increment": inc,
        "block_content": content,
        "nonce": int.from_bytes(raw[-4:], "little"),
        "digest": hashlib.sha256(raw).hexdigest(),
        "posted_txs": out["txs"],
        "posted_block_no": out["block_no"],
        "node_accepts": verdict,
    }


def mine_block_vectors():
    vecs = []
    for d in (4.0, 4, 5.3, 6.3):
        vecs.append(_vector(f"T0-v1 d={d!r}", O.T0_ADDRESS_V1, O.t0_mining_params(d), O.T0_TIMESTAMP))
    for d in (4.0, 5.3):
        vecs.append(_vector(f"T0-v2 d={d!r}", O.T0_ADDRESS_V2, O.t0_mining_params(d), O.T0_TIMESTAMP))
    # residue-class striding, the worker layout of miner.py:139-148
    for k, w in ((3, 8), (7, 8), (1, 3), (0, 5)):
        vecs.append(_vector(f"T0-v1 d=4.3 worker {k}/{w}", O.T0_ADDRESS_V1, O.t0_mining_params(4.3),
                            O.T0_TIMESTAMP, k, w))
    # genesis template: no 'hash'/'id' in last_block (miner.py:37-42)
    vecs.append(_vector("genesis d=3.0", O.T0_ADDRESS_V1,
                        {"difficulty": 3.0, "last_block": {}, "pending_transactions_hashes": []},
                        O.T0_TIMESTAMP))
    rng = random.Random(20261106)
    diffs = [1.0, 1.5, 2.0, 2.9, 3.0, 3.1, 3.7, 4.0, 4.2, 4.5, 4.6, 4.8, 4.9, 5.0, 5.1, 2.3, 3.4, 1.2]
    for i, d in enumerate(diffs):
        prev = rng.randbytes(32).hex()
        v2 = (i % 3 == 1)
        wallet = O.b58encode(bytes([42 + (i & 1)]) + rng.randbytes(32)) if v2 else rng.randbytes(64).hex()
        txs = sorted(rng.randbytes(32).hex() for _ in range(rng.randrange(0, 6)))
        ts = rng.randrange(1_600_000_000, 1_900_000_000)
        params = {"difficulty": d, "last_block": {"hash": prev, "id": rng.randrange(1, 10 ** 6)},
                  "pending_transactions_hashes": txs}
        vecs.append(_vector(f"random#{i} {'v2' if v2 else 'v1'} d={d}", wallet, params, ts))
    return vecs


def verifier_vectors(vecs):
    """check_block_is_valid verdicts for golden blocks at neighbouring difficulties and nonces."""
    out = []
    for v in vecs[:6]:
        lb = v["mining_params"]["last_block"]
        raw = bytes.fromhex(v["block_content"])
        for d in (v["mining_params"]["difficulty"], 3.9, 4.1, 4.5, 5.0, 5.2, 5.4, 5.9, 6.0, 6.2, 6.4, 6.9, 7.0):
            for dn in (0, 1):
                n = (v["nonce"] + dn) & 0xFFFFFFFF
                content = (raw[:-4] + n.to_bytes(4, "little")).hex()
                out.append({"block_content": content, "difficulty": d, "last_block": lb,
                            "valid": ref_loader.check_block_is_valid(content, d, lb)})
    return out


def merkle_vectors():
    """get_transactions_merkle_tree (upow/manager.py:365-378) run verbatim on random transaction blobs."""
    mgr = ref_loader.load_manager()
    rng = random.Random(3)
    cases = []
    for n in (0, 1, 2, 7, 40):
        txs = [rng.randbytes(rng.randrange(120, 400)).hex() for _ in range(n)]
        cases.append({"txs": txs, "merkle_tree": mgr.get_transactions_merkle_tree(txs)})
    return cases


def sweep_fixture():
    prefix = O.t0_prefix(5.0)
    rule = O.Rule(5.0, O.T0_PREV_HASH)
    # the C oracle must agree with the hashlib loop before it is trusted with 2^32 nonces
    lo, hi = 0, 1 << 22
    assert c_oracle.all_valid_nonces(prefix, rule, lo, hi)[0] == O.all_valid_nonces(prefix, rule, lo, hi)
    nonces, hashes = c_oracle.all_valid_nonces(prefix, rule, 0, 1 << 32)
    assert hashes == 1 << 32
    fp = O.nonce_set_hash(nonces)
    if fp != SURVEY_A3_HASH or len(nonces) != SURVEY_A3_COUNT:
        raise SystemExit(f"sweep mismatch vs SURVEY A.3: {len(nonces)} nonces, {fp}")
    with open(os.path.join(GOLDEN, "sweep_t0v1_d5.u32"), "wb") as f:
        f.write(struct.pack(f"<{len(nonces)}I", *nonces))
    meta = {"template": "T0-v1", "difficulty": 5.0, "prefix": prefix.hex(), "range": [0, 1 << 32],
            "count": len(nonces), "sha256_sorted_u32le": fp,
            "cumulative_per_2^28": [sum(1 for n in nonces if n < (k << 28)) for k in range(1, 17)],
            "generator": "oracle/upow_oracle.c full-message SHA-256, fingerprint equals SURVEY A.3 (hashlib loop)"}
    with open(os.path.join(GOLDEN, "sweep_t0v1_d5.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("sweep fixture:", len(nonces), fp)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    args = ap.parse_args()
    if not ref_loader.reference_available():
        raise SystemExit("reference tree not available; fixtures can only be generated in the build container")
    os.makedirs(GOLDEN, exist_ok=True)
    vecs = mine_block_vectors()
    with open(os.path.join(GOLDEN, "mine_block_vectors.json"), "w") as f:
        json.dump(vecs, f, indent=1)
    print(f"{len(vecs)} mine_block vectors; node accepts all: {all(v['node_accepts'] for v in vecs)}")
    ver = verifier_vectors(vecs)
    with open(os.path.join(GOLDEN, "verifier_vectors.json"), "w") as f:
        json.dump(ver, f, indent=1)
    print(f"{len(ver)} verifier vectors ({sum(v['valid'] for v in ver)} valid)")
    with open(os.path.join(GOLDEN, "merkle_tree_vectors.json"), "w") as f:
        json.dump(merkle_vectors(), f)
    if args.sweep:
        sweep_fixture()


if __name__ == "__main__":
    main()
