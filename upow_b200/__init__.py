"""upow_b200 -- B200-native nonce search for uPow (the miner.py hot path only).

``upow_b200.miner`` mirrors the reference's miner.py call surface; ``upow_b200.Context``
is the thin ctypes view of the C-ABI in include/upow_b200.h.  Nothing in this package
imports torch or the oracle, and nothing falls back to the CPU.
"""
from ._lib import (UPOW_CANCELLED, UPOW_EXHAUSTED, UPOW_OK, Context, UpowError, load, library_path, microbench,
                   selftest)
from .template import DifficultyRule, build_block_prefix, calculate_merkle_root, string_to_bytes

__all__ = ["Context", "UpowError", "load", "library_path", "microbench", "selftest", "DifficultyRule",
           "build_block_prefix", "calculate_merkle_root", "string_to_bytes", "UPOW_OK", "UPOW_EXHAUSTED",
           "UPOW_CANCELLED"]
