"""How the (timestamp, nonce) search space is split over GPUs / ranks.

The reference shards by residue class over OS processes with no communication
(miner.py:139-148).  Here the unit handed to one GPU is a whole nonce space of one timestamp
(or a contiguous tile of it); units are numbered in the search order "timestamp, then nonce", and
unit u of a step belongs to rank u // n_local, local GPU u % n_local.  No data-path collective:
ranks only meet in a barrier and a max-reduce of the elapsed time.
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


def reduce_elapsed(elapsed_local: float, dist=None) -> float:
    """Max over ranks of the elapsed time (the number the throughput is computed from)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return elapsed_local
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([elapsed_local], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def reduce_sum(x: float, dist=None) -> float:
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return x
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
