"""Block-template side of the miner, in Python exactly as the reference keeps it.

Everything here is O(1) per template and stays on the host:
the difficulty rule (miner.py:34-35,43-49), the prefix bytes (miner.py:63,74-82), the merkle
root (miner.py:15-18) and address decoding (upow/helpers.py:183-188).  The float expressions
are the reference's own (``difficulty % 1``, ``int(difficulty * 10)``, ``ceil(16 * (1 - frac))``)
so that no rounding difference can creep in.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass
from math import ceil

ENDIAN = "little"  # upow/constants.py:3
GENESIS_PREV_HASH = (18_884_643).to_bytes(32, ENDIAN).hex()  # miner.py:37-41
_B58 = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"


def b58decode(text: str) -> bytes:
    """Stand-in for base58.b58decode (33-byte compressed addresses only need this)."""
    value = 0
    for ch in text:
        value = value * 58 + _B58.index(ch)
    body = value.to_bytes((value.bit_length() + 7) // 8, "big")
    return b"\x00" * (len(text) - len(text.lstrip("1"))) + body


def string_to_bytes(text: str) -> bytes:
    """Hex first, base58 when that fails -- upow/helpers.py:183-188."""
    try:
        return bytes.fromhex(text)
    except ValueError:
        return b58decode(text)


def calculate_merkle_root(transactions) -> str:
    """sha256 over the concatenated tx hashes, miner.py:15-18."""
    joined = b"".join(bytes.fromhex(tx) for tx in transactions)
    return hashlib.sha256(joined).hexdigest()


@dataclass(frozen=True)
class DifficultyRule:
    """What the device needs to know about `difficulty` and the previous hash."""
    n_int: int            # int(difficulty): how many trailing hex chars of prev hash the digest must start with
    charset_count: int    # 16, or ceil(16*(1-frac)) when the fraction is > 0
    difficulty_field: int  # int(difficulty*10), the u16 stored in the block
    prev_hash: bytes

    @classmethod
    def derive(cls, difficulty, prev_hash_hex: str) -> "DifficultyRule":
        frac = difficulty % 1                     # miner.py:35
        count = 16
        if frac > 0:                              # miner.py:46-48
            count = ceil(16 * (1 - frac))
        return cls(int(difficulty), count, int(difficulty * 10), bytes.fromhex(prev_hash_hex))

    @property
    def target_prefix(self) -> str:
        return self.prev_hash.hex()[-self.n_int:]  # miner.py:43, including the N == 0 quirk

    def accepts(self, digest_hex: str) -> bool:
        """Host-side re-statement used for logging/asserts only (never for searching)."""
        if not digest_hex.startswith(self.target_prefix):
            return False
        return self.charset_count == 16 or digest_hex[self.n_int] in "0123456789abcdef"[: self.charset_count]


def build_block_prefix(prev_hash_hex: str, wallet_bytes: bytes, merkle_root_hex: str, timestamp: int,
                       difficulty) -> bytes:
    """[0x02]? + prev_hash + address + merkle + ts(u32 LE) + int(difficulty*10)(u16 LE), miner.py:74-82."""
    parts = [
        bytes.fromhex(prev_hash_hex),
        wallet_bytes,
        bytes.fromhex(merkle_root_hex),
        timestamp.to_bytes(4, byteorder=ENDIAN),
        int(difficulty * 10).to_bytes(2, ENDIAN),
    ]
    if len(wallet_bytes) == 33:
        parts.insert(0, (2).to_bytes(1, ENDIAN))
    return b"".join(parts)


def with_timestamp(prefix: bytes, timestamp: int) -> bytes:
    cut = len(prefix) - 6
    return prefix[:cut] + (timestamp & 0xFFFFFFFF).to_bytes(4, ENDIAN) + prefix[cut + 4:]
