"""How the (timestamp, nonce) search space is split over GPUs / ranks.

The reference shards by residue class over OS processes with no communication
(miner.py:139-148).  Here the unit handed to one GPU is a whole nonce space of one timestamp
(or a contiguous tile of it); units are numbered in the search order "timestamp, then nonce", and
unit u of a step belongs to rank u // n_local, local GPU u % n_local (the independent-ranks leg of bench.py).
Inside ONE library context the GPUs instead share every nonce space by lgroup interleave (lgroup_class below).
No data-path collective either way: ranks only meet in a barrier and a max-reduce of the elapsed time (bench.py).
"""
from __future__ import annotations

from typing import List


def unit_index(step: int, rank: int, local: int, world: int, n_local: int) -> int:
    """Global index (in search order) of the unit GPU `local` of `rank` works on in `step`."""
    return step * world * n_local + rank * n_local + local


def step_timestamps(step: int, rank: int, world: int, n_local: int, t0: int) -> List[int]:
    """Timestamps this rank's GPUs sweep in `step` (rollover rule: one timestamp per exhausted nonce space)."""
    return [t0 + unit_index(step, rank, j, world, n_local) for j in range(n_local)]


def split_range(lo: int, hi: int, parts: int, align: int = 1 << 16):
    """Contiguous shares of [lo, hi) with inner cuts aligned to `align` (no H value evaluated twice)."""
    cuts = []
    for i in range(parts + 1):
        c = lo + (hi - lo) * i // parts
        if 0 < i < parts:
            c = (c + align - 1) // align * align
        cuts.append(min(max(c, lo), hi))
    return [(cuts[i], cuts[i + 1]) for i in range(parts)]


LGROUPS = 2048   # a nonce is (H << 16) | L; an lgroup is 32 consecutive L values, the unit a warp owns per H


def lgroup_class(device_index: int, n_devices: int):
    """How ONE nonce space is shared by the GPUs of a context: device g searches the lgroups g, g + n, g + 2n, ...
    of every H, i.e. the nonces whose bits 5..15 are congruent to g mod n (upow_lib.cu contiguous_args).  All GPUs
    then walk H together and the lowest valid nonce is the minimum of the per-GPU minima."""
    return range(device_index, LGROUPS, n_devices)


def device_of_nonce(nonce: int, n_devices: int) -> int:
    return ((nonce & 0xFFFF) >> 5) % n_devices
