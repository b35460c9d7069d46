"""Block wire format on either side of the batch verifier: what chain sync does to a block
before and after hashing it.

Mirrors of the reference (file:line into /root/reference/upow/manager.py):
  * ``block_to_bytes(last_block_hash, block)``   :384-399  block dict -> the 138 / 108 content bytes
  * ``split_block_content(block_content_hex)``   :401-419  content -> (previous_hash, address, merkle_tree,
    timestamp, difficulty, random)
and, on top of ``upow_verify_blocks``, ``verify_chain_segment``: the hash / proof-of-work / linkage checks that
``create_blocks`` (upow/node/main.py:95-150) + ``check_block`` (upow/manager.py:422-460) apply to a run of
synced blocks one at a time, for the whole run in one device call.  Transactions, signatures and the database
are out of scope.

One deliberate difference: the reference turns the address bytes into a curve point and back
(``bytes_to_string`` -> fastecdsa) before returning them.  For every address that IS a valid point encoding that
round trip is the identity on the string, so this module formats the raw bytes directly (128 hex characters, or
base58 for the 33-byte form) and needs no elliptic-curve code.
"""
from __future__ import annotations

from decimal import Decimal
from io import BytesIO
from typing import List, Optional, Sequence, Tuple

from .template import ENDIAN, GENESIS_PREV_HASH, DifficultyRule, string_to_bytes

_B58 = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"


def _b58encode(b: bytes) -> str:
    n = int.from_bytes(b, "big")
    out = ""
    while n:
        n, r = divmod(n, 58)
        out = _B58[r] + out
    return "1" * (len(b) - len(b.lstrip(b"\0"))) + out


def block_to_bytes(last_block_hash: str, block: dict) -> bytes:
    """manager.py:384-399.  `block` needs address, merkle_tree, timestamp, difficulty, random."""
    address_bytes = string_to_bytes(block["address"])
    version = bytes([])
    if len(address_bytes) != 64:
        version = bytes([2])
    return (version
            + bytes.fromhex(last_block_hash)
            + address_bytes
            + bytes.fromhex(block["merkle_tree"])
            + block["timestamp"].to_bytes(4, byteorder=ENDIAN)
            + int(float(block["difficulty"]) * 10).to_bytes(2, ENDIAN)
            + block["random"].to_bytes(4, ENDIAN))


def split_block_content(block_content: str) -> Tuple[str, str, str, int, Decimal, int]:
    """manager.py:401-419, same asserts (138 bytes = version 1; otherwise a version byte > 1, version 2 = 108 bytes)."""
    _bytes = bytes.fromhex(block_content)
    stream = BytesIO(_bytes)
    if len(_bytes) == 138:
        version = 1
    else:
        version = int.from_bytes(stream.read(1), ENDIAN)
        assert version > 1
        if version == 2:
            assert len(_bytes) == 108
        else:
            raise NotImplementedError()
    previous_hash = stream.read(32).hex()
    raw = stream.read(64 if version == 1 else 33)
    address = raw.hex() if version == 1 else _b58encode(raw)
    merkle_tree = stream.read(32).hex()
    timestamp = int.from_bytes(stream.read(4), ENDIAN)
    difficulty = int.from_bytes(stream.read(2), ENDIAN) / Decimal(10)
    random = int.from_bytes(stream.read(4), ENDIAN)
    return previous_hash, address, merkle_tree, timestamp, difficulty, random


def verify_chain_segment(ctx, last_block: dict, blocks: Sequence[dict]) -> Tuple[Optional[int], List[str]]:
    """A run of consecutive blocks as chain sync receives them (each a dict with id, hash, address, merkle_tree,
    timestamp, difficulty, random and optionally content), to be appended after `last_block` ({} at genesis).

    Per block, what the node checks about the header while syncing:
      * content = block["content"] or block_to_bytes(previous block's hash, block)      node/main.py:118-120
      * sha256(content) == block["hash"]                                                 (how the stored hash is defined)
      * proof of work at the block's difficulty against the previous block's hash        manager.py:130-151
      * the content's previous-hash field is the previous block's hash                   manager.py:439-443
      * id follows, timestamp grows                                                      node/main.py:140, manager.py:445-455
    All SHA-256 work runs in ONE upow_verify_blocks call.  Returns (index of the first bad block or None, reasons).
    """
    prev_hash = last_block["hash"] if "hash" in last_block else GENESIS_PREV_HASH
    prev_id = last_block.get("id", 0)
    prev_ts = last_block.get("timestamp", 0)
    records, contents = [], []
    for b in blocks:
        content = b.get("content")
        if isinstance(content, str):
            content = bytes.fromhex(content)
        if not content:
            content = block_to_bytes(prev_hash, b)
        rule = DifficultyRule.derive(float(b["difficulty"]), prev_hash)
        records.append((content, rule.prev_hash, rule.n_int, rule.charset_count))
        contents.append(content)
        prev_hash = b["hash"]        # the next block claims to build on this one; the claim itself is checked below
    verdicts, digests, _ = ctx.verify_blocks(records)
    reasons: List[str] = []
    chain_hash = last_block["hash"] if "hash" in last_block else GENESIS_PREV_HASH
    genesis = "hash" not in last_block
    for i, b in enumerate(blocks):
        why = None
        first_ever = genesis and i == 0
        if verdicts[i] == 255:
            why = "content is neither 138 nor 108 bytes"
        elif digests[i].hex() != b["hash"]:
            why = "sha256(content) differs from the block's hash"
        elif verdicts[i] != 1 and not first_ever:        # manager.py:135-136: with no previous block anything is valid
            why = "block not valid"
        elif not first_ever and contents[i][(1 if len(contents[i]) == 108 else 0):][:32].hex() != chain_hash:
            why = "Previous hash is not matched"
        elif b.get("id", prev_id + 1) != prev_id + 1:
            why = "block id does not follow"
        elif b["timestamp"] <= prev_ts:
            why = "timestamp younger than previous block"
        if why:
            reasons.append(why)
            return i, reasons
        chain_hash, prev_id, prev_ts = b["hash"], prev_id + 1, b["timestamp"]
    return None, reasons
