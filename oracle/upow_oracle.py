"""CPU oracle for the uPow nonce search -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this module.  The product path
(``upow_b200/``) never does: it calls the CUDA library through the C-ABI and
fails loudly when that library is missing.

What this restates (reference = /root/reference, pure Python):

* rule derivation                miner.py:34-35,43-49
* validity predicate             miner.py:51-56 (fractional) / miner.py:60-61 (integer)
* prefix construction            miner.py:63,74-82
* nonce encoding                 miner.py:87   (u32 little-endian, upow/constants.py:3)
* merkle root                    miner.py:15-18
* address decoding               upow/helpers.py:183-188 (hex first, base58 on ValueError)
* node-side verifier             upow/manager.py:130-151 (Decimal arithmetic)

The SHA-256 itself lives in a third-party dependency that is not under
/root/reference: CPython ``hashlib`` -> OpenSSL EVP (stdlib, no version pin in
requirements.txt; README names Python 3.11.9).  SHA-256 is FIPS 180-4, so the
oracle calls the very same ``hashlib.sha256`` the reference calls.

Parity pinning: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so this oracle is pinned against outputs of the
reference itself -- ``oracle/gen_golden.py`` imports /root/reference/miner.py
and upow/manager.py verbatim (with stub modules for the uninstalled
third-party imports), runs ``mine_block`` / ``check_block_is_valid`` and
commits the results under ``tests/golden/``; ``tests/test_oracle.py`` checks
every function below against those fixtures and the FIPS 180-4 KATs.
"""
from __future__ import annotations

import hashlib
import struct
import time
from decimal import Decimal
from math import ceil, floor

ENDIAN = "little"  # upow/constants.py:3
HEX_CHARSET = "0123456789abcdef"  # miner.py:45
GENESIS_PREV_HASH = (18_884_643).to_bytes(32, ENDIAN).hex()  # miner.py:37-41

_B58_ALPHABET = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz"


# --------------------------------------------------------------------------
# helpers the reference takes from third-party packages
# --------------------------------------------------------------------------
def b58decode(s: str) -> bytes:
    """base58 (bitcoin alphabet) decode; stands in for ``base58.b58decode``."""
    n = 0
    for ch in s:
        n = n * 58 + _B58_ALPHABET.index(ch)
    body = n.to_bytes((n.bit_length() + 7) // 8, "big") if n else b""
    pad = len(s) - len(s.lstrip("1"))
    return b"\x00" * pad + body


def b58encode(b: bytes) -> str:
    n = int.from_bytes(b, "big")
    out = ""
    while n:
        n, r = divmod(n, 58)
        out = _B58_ALPHABET[r] + out
    pad = len(b) - len(b.lstrip(b"\x00"))
    return "1" * pad + out


def string_to_bytes(string: str) -> bytes:
    """upow/helpers.py:183-188."""
    try:
        return bytes.fromhex(string)
    except ValueError:
        return b58decode(string)


def calculate_merkle_root(transactions) -> str:
    """miner.py:15-18."""
    return hashlib.sha256(b"".join(bytes.fromhex(tx) for tx in transactions)).hexdigest()


def transactions_merkle_tree(txs_hex) -> str:
    """upow/manager.py:365-378: sort the transactions' bytes, sha256 each, sha256 the concatenation."""
    ordered = sorted(bytes.fromhex(t) for t in txs_hex)
    return hashlib.sha256(b"".join(hashlib.sha256(b).digest() for b in ordered)).hexdigest()


# --------------------------------------------------------------------------
# rule + prefix (miner.py:34-49, 74-82)
# --------------------------------------------------------------------------
class Rule:
    """Difficulty rule exactly as ``mine_block`` derives it."""

    __slots__ = ("difficulty", "target_prefix", "allowed_chars", "int_difficulty", "fractional")

    def __init__(self, difficulty, prev_hash_hex: str):
        self.difficulty = difficulty
        self.fractional = difficulty % 1  # miner.py:35
        self.target_prefix = prev_hash_hex[-int(difficulty):]  # miner.py:43 (N==0 -> whole string)
        self.allowed_chars = HEX_CHARSET
        self.int_difficulty = int(difficulty)
        if self.fractional > 0:  # miner.py:46-49
            self.allowed_chars = HEX_CHARSET[: ceil(16 * (1 - self.fractional))]

    @property
    def charset_count(self) -> int:
        return len(self.allowed_chars)

    def is_valid_hex(self, digest_hex: str) -> bool:
        if self.fractional > 0:  # miner.py:51-56
            return (digest_hex.startswith(self.target_prefix)
                    and digest_hex[self.int_difficulty] in self.allowed_chars)
        return digest_hex.startswith(self.target_prefix)  # miner.py:60-61

    def is_valid_block(self, block_data: bytes) -> bool:
        return self.is_valid_hex(hashlib.sha256(block_data).hexdigest())


def build_prefix(prev_hash_hex: str, wallet_bytes: bytes, merkle_root_hex: str,
                 timestamp: int, difficulty) -> bytes:
    """miner.py:74-82."""
    block_prefix = (
        bytes.fromhex(prev_hash_hex)
        + wallet_bytes
        + bytes.fromhex(merkle_root_hex)
        + timestamp.to_bytes(4, byteorder=ENDIAN)
        + int(difficulty * 10).to_bytes(2, ENDIAN)
    )
    if len(wallet_bytes) == 33:
        block_prefix = (2).to_bytes(1, ENDIAN) + block_prefix
    return block_prefix


def set_timestamp(prefix: bytes, timestamp: int) -> bytes:
    """Same prefix with another timestamp (the ts+1 rollover rule, SURVEY 8(d) cfg 4)."""
    off = len(prefix) - 6
    return prefix[:off] + (timestamp & 0xFFFFFFFF).to_bytes(4, ENDIAN) + prefix[off + 4:]


def get_timestamp(prefix: bytes) -> int:
    off = len(prefix) - 6
    return int.from_bytes(prefix[off:off + 4], ENDIAN)


def node_check_block_is_valid(block_content_hex: str, difficulty, last_block: dict) -> bool:
    """upow/manager.py:130-151 restated (Decimal arithmetic, node side)."""
    block_hash = hashlib.sha256(bytes.fromhex(block_content_hex)).hexdigest()
    if "hash" not in last_block:
        return True
    last_block_hash = last_block["hash"]
    difficulty = Decimal(str(difficulty))
    decimal = difficulty % 1
    n = floor(difficulty)
    if decimal > 0:
        count = ceil(16 * (1 - decimal))
        return block_hash.startswith(last_block_hash[-n:]) and block_hash[n] in HEX_CHARSET[:count]
    return block_hash.startswith(last_block_hash[-n:])


# --------------------------------------------------------------------------
# the hot loop (miner.py:83-98) and the searches built on it
# --------------------------------------------------------------------------
def min_valid_nonce(prefix: bytes, rule: Rule, start_nonce: int = 0, nonce_increment: int = 1,
                    max_nonce: int = 1 << 32):
    """First nonce in start, start+inc, ... < max_nonce whose candidate is valid, else None.

    The reference raises OverflowError at nonce >= 2**32 (miner.py:87); returning
    None is this oracle's spelling of "range exhausted".
    """
    sha = hashlib.sha256
    nonce = start_nonce
    valid = rule.is_valid_hex
    while nonce < max_nonce:
        if valid(sha(prefix + nonce.to_bytes(4, ENDIAN)).hexdigest()):
            return nonce
        nonce += nonce_increment
    return None


def miner_loop(prefix: bytes, rule: Rule, start_nonce: int = 0, nonce_increment: int = 1,
               max_nonce: int = 1 << 32, progress=None):
    """The hot loop with the SHAPE of miner.py:83-98, statement for statement: candidate = prefix + nonce LE32,
    predicate on the hex digest (miner.py:51-56 / 60-61), then the increment and the per-iteration
    `% check_interval` progress test (miner.py:89-95; the print goes to `progress` if given, the 90 s staleness
    test is left to the caller).  Same answers as min_valid_nonce; this is the form bench.py times as the CPU arm,
    because it pays every per-iteration cost the reference pays.  Returns None where the reference would raise
    OverflowError (nonce >= 2**32) or when max_nonce is reached."""
    sha = hashlib.sha256
    target_prefix = rule.target_prefix
    allowed_chars = rule.allowed_chars if rule.fractional > 0 else None
    idiff = rule.int_difficulty
    if allowed_chars is not None:
        def is_valid_block(block_data: bytes) -> bool:
            block_hash = sha(block_data).hexdigest()
            return block_hash.startswith(target_prefix) and block_hash[idiff] in allowed_chars
    else:
        def is_valid_block(block_data: bytes) -> bool:
            return sha(block_data).hexdigest().startswith(target_prefix)
    nonce = start_nonce
    check_interval = 5000000 * nonce_increment
    t0 = time.time()
    while nonce < max_nonce:
        if is_valid_block(prefix + nonce.to_bytes(4, ENDIAN)):
            return nonce
        nonce += nonce_increment
        if (nonce - start_nonce) % check_interval == 0:
            elapsed = time.time() - t0
            if progress is not None:
                progress(nonce, elapsed)
    return None


def all_valid_nonces(prefix: bytes, rule: Rule, lo: int, hi: int):
    """Sorted list of every valid nonce in [lo, hi)."""
    sha = hashlib.sha256
    valid = rule.is_valid_hex
    return [n for n in range(lo, hi) if valid(sha(prefix + n.to_bytes(4, ENDIAN)).hexdigest())]


def first_valid_with_rollover(prefix: bytes, rule: Rule, max_rollovers: int,
                              max_nonce: int = 1 << 32):
    """Lexicographically first (timestamp, nonce) under the rule ts<-ts+1, nonce<-0."""
    ts0 = get_timestamp(prefix)
    for r in range(max_rollovers + 1):
        p = set_timestamp(prefix, ts0 + r)
        n = min_valid_nonce(p, rule, 0, 1, max_nonce)
        if n is not None:
            return ts0 + r, n
    return None


def digest_of(prefix: bytes, nonce: int) -> bytes:
    return hashlib.sha256(prefix + nonce.to_bytes(4, ENDIAN)).digest()


def nonce_set_hash(nonces) -> str:
    """SURVEY Appendix A.3 fingerprint of a set of nonces."""
    return hashlib.sha256(b"".join(n.to_bytes(4, ENDIAN) for n in sorted(nonces))).hexdigest()


# --------------------------------------------------------------------------
# pure-Python SHA-256 compression: used to check the midstate / tail-block
# layout the CUDA library relies on.  Small cases only.
# --------------------------------------------------------------------------
_K = [
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2,
]
_H0 = [0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19]
_M = 0xFFFFFFFF


def _rotr(x, n):
    return ((x >> n) | (x << (32 - n))) & _M


def sha256_compress(state, block: bytes):
    """One FIPS 180-4 compression; state = list of 8 u32, block = 64 bytes."""
    w = list(struct.unpack(">16I", block))
    for t in range(16, 64):
        s0 = _rotr(w[t - 15], 7) ^ _rotr(w[t - 15], 18) ^ (w[t - 15] >> 3)
        s1 = _rotr(w[t - 2], 17) ^ _rotr(w[t - 2], 19) ^ (w[t - 2] >> 10)
        w.append((w[t - 16] + s0 + w[t - 7] + s1) & _M)
    a, b, c, d, e, f, g, h = state
    for t in range(64):
        t1 = (h + (_rotr(e, 6) ^ _rotr(e, 11) ^ _rotr(e, 25)) + ((e & f) ^ (~e & g & _M)) + _K[t] + w[t]) & _M
        t2 = ((_rotr(a, 2) ^ _rotr(a, 13) ^ _rotr(a, 22)) + ((a & b) ^ (a & c) ^ (b & c))) & _M
        h, g, f, e, d, c, b, a = g, f, e, (d + t1) & _M, c, b, a, (t1 + t2) & _M
    return [(x + y) & _M for x, y in zip(state, (a, b, c, d, e, f, g, h))]


def midstate_and_tail(prefix: bytes):
    """(state after the constant leading blocks, 64-byte tail block with nonce=0).

    138-byte messages: 2 constant blocks, tail = bytes 128..137 + padding;
    108-byte messages: 1 constant block, tail = bytes 64..107 + padding.
    """
    msg_len = len(prefix) + 4
    assert msg_len in (138, 108)
    n_const = (msg_len // 64)
    state = list(_H0)
    for i in range(n_const):
        state = sha256_compress(state, prefix[64 * i:64 * i + 64])
    rest = prefix[64 * n_const:] + b"\x00\x00\x00\x00"
    tail = rest + b"\x80" + b"\x00" * (64 - len(rest) - 1 - 8) + struct.pack(">Q", msg_len * 8)
    assert len(tail) == 64
    return state, tail


def tail_digest(prefix: bytes, nonce: int) -> bytes:
    """SHA-256 via midstate + one tail compression (must equal hashlib)."""
    state, tail = midstate_and_tail(prefix)
    off = len(prefix) - 64 * ((len(prefix) + 4) // 64)
    tail = tail[:off] + nonce.to_bytes(4, ENDIAN) + tail[off + 4:]
    return struct.pack(">8I", *sha256_compress(state, tail))


# --------------------------------------------------------------------------
# SURVEY 8(d) template T0 (deterministic synthetic template used everywhere)
# --------------------------------------------------------------------------
def _h(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()


T0_PREV_HASH = _h(b"prev-block")
T0_ADDRESS_V1 = _h(b"addr-x") + _h(b"addr-y")
T0_ADDRESS_V2 = b58encode(b"\x2a" + hashlib.sha256(b"addr-c").digest())
T0_TX_HASHES = [_h(b"tx0"), _h(b"tx1"), _h(b"tx2")]
T0_TIMESTAMP = 1_700_000_000
T0_LAST_BLOCK_ID = 1000


def t0_mining_params(difficulty):
    return {
        "difficulty": difficulty,
        "last_block": {"hash": T0_PREV_HASH, "id": T0_LAST_BLOCK_ID},
        "pending_transactions_hashes": list(T0_TX_HASHES),
    }


def t0_prefix(difficulty, version: int = 1, timestamp: int = T0_TIMESTAMP) -> bytes:
    addr = T0_ADDRESS_V1 if version == 1 else T0_ADDRESS_V2
    return build_prefix(T0_PREV_HASH, string_to_bytes(addr), calculate_merkle_root(T0_TX_HASHES),
                        timestamp, difficulty)


def batch_template(i: int, difficulty, timestamp: int = T0_TIMESTAMP):
    """cfg 5 template i: per-template address and tx list."""
    le = i.to_bytes(4, ENDIAN)
    addr = _h(b"addr-x" + le) + _h(b"addr-y" + le)
    txs = [_h(b"tx" + le + bytes([j])) for j in range(3)]
    return build_prefix(T0_PREV_HASH, string_to_bytes(addr), calculate_merkle_root(txs),
                        timestamp, difficulty)
