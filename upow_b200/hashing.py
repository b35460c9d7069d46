"""Transaction-hash and merkle-root helpers on top of upow_sha256_batch.

Mirrors, with the per-transaction hashing done in one GPU call:
  * get_mining_info's  pending_transactions_hashes = [sha256(tx) for tx in sorted(pending)]   upow/node/main.py:681,692
  * get_transactions_merkle_tree(transactions)                                             upow/manager.py:365-378
  * Transaction.hash = sha256(bytes.fromhex(tx.hex()))                                   upow/upow_transactions/transaction.py:85-88
Only useful when many transactions are hashed at once (a block holds up to ~8300, README.md:16; the
template multiplier re-roots many subsets).
"""
from __future__ import annotations

from typing import List, Sequence


def transaction_hashes(ctx, txs_hex: Sequence[str]) -> List[str]:
    """sha256 of every transaction (hex in, hex digest out), input order kept."""
    digests, _ = ctx.sha256_many(bytes.fromhex(t) for t in txs_hex)
    return [d.hex() for d in digests]


def pending_transactions_hashes(ctx, pending_txs_hex: Sequence[str]) -> List[str]:
    """What get_mining_info returns: hashes of the pending transactions in sorted-hex order."""
    return transaction_hashes(ctx, sorted(pending_txs_hex))


def transactions_merkle_tree(ctx, txs_hex: Sequence[str]) -> str:
    """upow/manager.py:365-378: sort the transactions' bytes, hash each, hash the concatenation."""
    ordered = sorted(bytes.fromhex(t) for t in txs_hex)
    digests, _ = ctx.sha256_many(ordered)
    root, _ = ctx.sha256_many([b"".join(digests)])
    return root[0].hex()
