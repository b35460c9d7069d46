"""Generate tests/golden/* from the reference itself -- TEST INFRASTRUCTURE ONLY.

Run in the build container (needs /root/reference):

    python -m oracle.gen_golden            # mine_block / verifier vectors (about a minute)
    python -m oracle.gen_golden --sweep    # + the 2^32 sweep fixture (minutes, C oracle)
    python -m oracle.gen_golden --cfg4     # only: minimality pin of BASELINE config 4 at difficulty 9.0 (C oracle, ~3e9 hashes)
    python -m oracle.gen_golden --cfg5     # only: minimum nonce of all 4096 config-5 templates at 6.0 (C oracle, ~7e10 hashes)

* ``mine_block_vectors.json``: every entry is ``miner.mine_block(start, inc, params)``
  (miner.py:30-114) run VERBATIM with only ``requests.post`` and ``timestamp()``
  replaced, plus the verdict of ``upow.manager.check_block_is_valid``
  (upow/manager.py:130-151) on the resulting block, also run verbatim.
* ``verifier_vectors.json``: check_block_is_valid verdicts on near-miss blocks.
* ``merkle_tree_vectors.json``: get_transactions_merkle_tree (upow/manager.py:365-378) on random blobs.
* ``cfg4_t0v1_d9.json``: the first valid nonce of T0-v1 at difficulty 9.0 in the order of miner.py:86-89
  (timestamp 1 700 000 000, nonces 0, 1, 2, ...), found by the C oracle walking every nonce below it.
* ``cfg5_min_nonces.u32`` (+ ``.json``): the first valid nonce of each of the 4096 config-5 templates at
  difficulty 6.0, one sequential C-oracle walk from nonce 0 per template.
  Both are preceded by a check of the C oracle against the plain hashlib loop on 2^28 nonces.
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
import time

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


def _hashlib_slice(args):
    prefix, difficulty, lo, hi = args
    return O.all_valid_nonces(prefix, O.Rule(difficulty, O.T0_PREV_HASH), lo, hi)


def check_c_oracle_against_hashlib(log, workers):
    """SURVEY 8(c) tier T2: before the C oracle is trusted with sizes Python cannot reach, its valid-nonce set
    over 2^28 nonces must equal the plain hashlib loop's (miner.py:51-61 restated in upow_oracle.py)."""
    import multiprocessing as mp
    prefix, d = O.t0_prefix(5.0), 5.0
    n, parts = 1 << 28, 64
    t0 = time.time()
    with mp.get_context("fork").Pool(workers) as pool:
        py = sum(pool.map(_hashlib_slice, [(prefix, d, n * i // parts, n * (i + 1) // parts) for i in range(parts)]), [])
    t1 = time.time()
    c, hashes = c_oracle.all_valid_nonces(prefix, O.Rule(d, O.T0_PREV_HASH), 0, n, threads=workers)
    assert hashes == n and c == py, "C oracle disagrees with the hashlib loop"
    log(f"C oracle == hashlib loop on [0, 2^28) of T0-v1 at 5.0: {len(py)} valid nonces "
        f"(hashlib {t1 - t0:.0f} s, C {time.time() - t1:.0f} s, {workers} workers)")


def cfg4_fixture(log, workers):
    """BASELINE config 4 at difficulty 9.0: the first valid (timestamp, nonce) in the order of miner.py:86-89."""
    d = 9.0
    prefix, rule = O.t0_prefix(d), O.Rule(d, O.T0_PREV_HASH)
    t0 = time.time()
    nonce, hashes = c_oracle.min_valid_nonce(prefix, rule, 0, 1, 1 << 32, threads=workers)
    assert nonce is not None, "no valid nonce at timestamp T0: rollover needed, extend this generator"
    block = prefix + nonce.to_bytes(4, "little")
    digest = hashlib.sha256(block).hexdigest()
    assert rule.is_valid_block(block) and O.node_check_block_is_valid(block.hex(), d, {"hash": O.T0_PREV_HASH})
    meta = {"template": "T0-v1", "difficulty": d, "timestamp": O.T0_TIMESTAMP, "rollovers": 0, "nonce": nonce,
            "digest": digest, "hashes_walked": hashes,
            "generator": "oracle/upow_oracle.c: every nonce below the answer hashed (full message) and rejected"}
    with open(os.path.join(GOLDEN, "cfg4_t0v1_d9.json"), "w") as f:
        json.dump(meta, f, indent=1)
    log(f"cfg4 9.0: first valid nonce {nonce} digest {digest} after {hashes} hashes, {time.time() - t0:.0f} s")


def _cfg5_one(i):
    d = 6.0
    n, hashes = c_oracle.min_valid_nonce(O.batch_template(i, d), O.Rule(d, O.T0_PREV_HASH), 0, 1, 1 << 32, threads=1)
    return i, (-1 if n is None else n), hashes


def cfg5_fixture(log, workers):
    """BASELINE config 5: minimum valid nonce of each of the 4096 templates at difficulty 6.0."""
    import multiprocessing as mp
    t0 = time.time()
    mins, total = [0] * 4096, 0
    with mp.get_context("fork").Pool(workers) as pool:
        for k, (i, n, hashes) in enumerate(pool.imap_unordered(_cfg5_one, range(4096), chunksize=4)):
            assert n >= 0, f"template {i} has no valid nonce below 2^32"
            mins[i] = n
            total += hashes
            if (k + 1) % 256 == 0:
                log(f"cfg5: {k + 1}/4096 templates, {total} hashes, {time.time() - t0:.0f} s")
    raw = struct.pack("<4096I", *mins)
    fp = hashlib.sha256(raw).hexdigest()
    with open(os.path.join(GOLDEN, "cfg5_min_nonces.u32"), "wb") as f:
        f.write(raw)
    meta = {"templates": "oracle.upow_oracle.batch_template(i, 6.0), i = 0..4095 (SURVEY 8(d) cfg 5)", "difficulty": 6.0,
            "count": 4096, "sha256_u32le_in_template_order": fp, "sum_min_nonce_plus_1": sum(mins) + 4096,
            "hashes_walked": total, "max_min_nonce": max(mins),
            "generator": "oracle/upow_oracle.c, one sequential walk from nonce 0 per template (full-message SHA-256)"}
    with open(os.path.join(GOLDEN, "cfg5_min_nonces.json"), "w") as f:
        json.dump(meta, f, indent=1)
    log(f"cfg5 fixture: fingerprint {fp}, {total} hashes, {time.time() - t0:.0f} s")


def blockfmt_vectors():
    """block_to_bytes / split_block_content (upow/manager.py:384-419) run verbatim on random blocks.  The reference's
    split_block_content sends the address through fastecdsa (bytes -> curve point -> string), which is not
    installed here and is the identity for valid points; for this run bytes_to_string is replaced by the plain
    formatting of the bytes (hex / base58), everything else is the reference's own code."""
    mgr = ref_loader.load_manager()
    mgr.bytes_to_string = lambda b: b.hex() if len(b) == 64 else O.b58encode(b)
    rng = random.Random(20261107)
    out = []
    for i in range(24):
        v2 = i % 3 == 1
        address = O.b58encode(bytes([42 + (i & 1)]) + rng.randbytes(32)) if v2 else rng.randbytes(64).hex()
        d = rng.choice(["6.0", "6.3", "7.9", "10.5", "0.1", "25.5", 6, 6.5])
        block = {"address": address, "merkle_tree": rng.randbytes(32).hex(), "timestamp": rng.randrange(1, 1 << 32),
                 "difficulty": d, "random": rng.randrange(0, 1 << 32)}
        prev = rng.randbytes(32).hex()
        content = mgr.block_to_bytes(prev, block)
        split = mgr.split_block_content(content.hex())
        out.append({"last_block_hash": prev, "block": block, "content": content.hex(),
                    "split": [split[0], split[1], split[2], split[3], str(split[4]), split[5]]})
    with open(os.path.join(GOLDEN, "blockfmt_vectors.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(f"{len(out)} block wire-format vectors")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--blockfmt", action="store_true", help="only: block_to_bytes / split_block_content vectors")
    ap.add_argument("--cfg4", action="store_true")
    ap.add_argument("--cfg5", action="store_true")
    ap.add_argument("--workers", type=int, default=os.cpu_count() or 1)
    args = ap.parse_args()
    if args.blockfmt:
        blockfmt_vectors()
        return
    if args.cfg4 or args.cfg5:
        os.makedirs(GOLDEN, exist_ok=True)
        logf = open(os.path.join(GOLDEN, "cfg_pins.log"), "a")

        def log(msg):
            line = time.strftime("%Y-%m-%dT%H:%M:%SZ ", time.gmtime()) + msg
            print(line, flush=True)
            logf.write(line + "\n")
            logf.flush()
        check_c_oracle_against_hashlib(log, args.workers)
        if args.cfg4:
            cfg4_fixture(log, args.workers)
        if args.cfg5:
            cfg5_fixture(log, args.workers)
        return
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
