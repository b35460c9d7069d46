"""Template multiplier: many block templates from ONE get_mining_info reply.

Why: the protocol gives a template only 2^32 nonces, and a timestamp is valid for every second in
(last_block.timestamp, node's now] (upow/manager.py:445-460) -- about 4.3 GH/s of fresh search space
per wall-clock second, a quarter of one B200.  The node leaves two more degrees of freedom:
  * any subset of the pending transactions it knows may be confirmed (upow/node/main.py:580-600); the
    block's merkle root must equal sha256 over the sha256 of the chosen transactions in sorted-hex order
    (upow/manager.py:365-378) -- which is what miner.py:15-18 computes from the node's already sorted
    hash list, so every *order-preserving* subset of pending_transactions_hashes is a valid choice;
  * every timestamp of the window.
This module enumerates those templates and races them with upow_search_any.  The reference itself
never does this (its CPU rate is ~10^6 H/s); semantics of each produced block are exactly the
reference's: same prefix layout (miner.py:74-82), same rule (miner.py:43-56), same POST body.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterator, List, Optional, Sequence

from .template import (ENDIAN, GENESIS_PREV_HASH, DifficultyRule, build_block_prefix, calculate_merkle_root,
                       string_to_bytes)


@dataclass(frozen=True)
class Candidate:
    prefix: bytes
    timestamp: int
    tx_hashes: tuple       # the transactions this template confirms (order preserved)
    merkle_root: str


def tx_subsets(tx_hashes: Sequence[str], limit: int) -> Iterator[tuple]:
    """Order-preserving subsets, most transactions first: all, then each single omission, then each
    pair of omissions, ... (a miner prefers to confirm as much as possible)."""
    n = len(tx_hashes)
    produced = 0
    from itertools import combinations
    for dropped in range(0, n + 1):
        for omit in combinations(range(n - 1, -1, -1), dropped):
            skip = set(omit)
            yield tuple(h for i, h in enumerate(tx_hashes) if i not in skip)
            produced += 1
            if produced >= limit:
                return


def candidates(mining_info: dict, wallet: str, now: int, max_templates: int,
               max_tx_subsets: Optional[int] = None) -> List[Candidate]:
    """Up to max_templates distinct templates: newest timestamps first within each transaction subset."""
    difficulty = mining_info["difficulty"]
    last_block = mining_info["last_block"]
    prev_hash = last_block["hash"] if "hash" in last_block else GENESIS_PREV_HASH
    last_ts = last_block.get("timestamp", 0)
    if not isinstance(last_ts, int):
        last_ts = 0
    first_ts = max(last_ts + 1, now - 3600)            # never reach back further than an hour
    if first_ts > now:
        return []
    window = now - first_ts + 1
    tx_hashes = list(mining_info["pending_transactions_hashes"])
    assert all(len(tx) == 64 for tx in tx_hashes)     # miner.py:69
    wallet_bytes = string_to_bytes(wallet)
    n_subsets = max_tx_subsets or max(1, -(-max_templates // window))
    out: List[Candidate] = []
    for subset in tx_subsets(tx_hashes, n_subsets):
        merkle = calculate_merkle_root(subset)
        for ts in range(now, first_ts - 1, -1):
            out.append(Candidate(build_block_prefix(prev_hash, wallet_bytes, merkle, ts, difficulty), ts, subset, merkle))
            if len(out) >= max_templates:
                return out
    return out


def mine_any(ctx, mining_info: dict, wallet: str, now: int, max_templates: int = 4096, max_nonce: int = 0):
    """Race the candidate templates on the GPUs; returns None or a dict with everything push_block needs."""
    difficulty = mining_info["difficulty"]
    last_block = mining_info["last_block"]
    prev_hash = last_block["hash"] if "hash" in last_block else GENESIS_PREV_HASH
    rule = DifficultyRule.derive(difficulty, prev_hash)
    cands = candidates(mining_info, wallet, now, max_templates)
    if not cands:
        return None
    rc, idx, nonce, digest, stats = ctx.search_any(
        [(c.prefix, rule.prev_hash, rule.n_int, rule.charset_count) for c in cands], max_nonce)
    if rc != 0:
        return None
    c = cands[idx]
    block = c.prefix + nonce.to_bytes(4, ENDIAN)
    return {"block_content": block.hex(), "txs": list(c.tx_hashes), "block_no": last_block.get("id", 0) + 1,
            "timestamp": c.timestamp, "nonce": nonce, "digest": digest.hex(), "template_index": idx,
            "templates": len(cands), "hashes_done": stats.hashes_done, "elapsed_s": stats.elapsed_s}
