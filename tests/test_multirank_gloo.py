"""world_size-2 CPU tests (gloo) of the N>1 host logic: the rank -> (timestamp, nonce) partition is a
disjoint cover in search order, the timing reduction is a max over ranks, and the CPU arm of bench.py
prints exactly one line from rank 0."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT
from upow_b200 import sharding


def test_partition_is_a_disjoint_cover_in_search_order():
    t0 = 1_700_000_000
    for world, n_local in ((1, 1), (2, 1), (4, 1), (8, 1), (1, 8), (2, 4)):
        seen = []
        for step in range(5):
            for rank in range(world):
                seen += sharding.step_timestamps(step, rank, world, n_local, t0)
        assert sorted(seen) == list(range(t0, t0 + 5 * world * n_local)), (world, n_local)
        assert len(set(seen)) == len(seen)


def test_lgroup_interleave_is_a_disjoint_cover():
    """Inside one context: every nonce belongs to exactly one GPU's lgroup class, for any GPU count."""
    for n in (1, 2, 3, 4, 5, 8):
        classes = [set(sharding.lgroup_class(g, n)) for g in range(n)]
        assert set().union(*classes) == set(range(sharding.LGROUPS))
        assert sum(len(c) for c in classes) == sharding.LGROUPS
        assert max(len(c) for c in classes) - min(len(c) for c in classes) <= 1
        for nonce in (0, 31, 32, 65535, 164867, 4084426313, (1 << 32) - 1):
            g = sharding.device_of_nonce(nonce, n)
            assert ((nonce & 0xFFFF) >> 5) in classes[g]


def test_split_range_alignment():
    for lo, hi, parts in ((0, 1 << 32, 8), (5, 1 << 30, 3), (12345, 7_000_001, 2), (0, 100, 4), (0, 0, 2)):
        shares = sharding.split_range(lo, hi, parts)
        assert shares[0][0] == lo and shares[-1][1] == hi
        for (a, b), (c, d) in zip(shares, shares[1:]):
            assert b == c and a <= b
        for a, b in shares[:-1]:
            assert b % (1 << 16) == 0 or b == hi


_WORKER = r"""
import os, sys, json
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
from upow_b200 import sharding
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
mine = []
for step in range(3):
    mine += sharding.step_timestamps(step, rank, world, 1, 1000)
gathered = [None] * world
dist.all_gather_object(gathered, mine)
elapsed = bench.reduce_elapsed(1.0 + rank, dist)          # rank 1 is the slow one
total = bench.reduce_sum(float(len(mine)), dist)
if rank == 0:
    print(json.dumps({{"gathered": gathered, "elapsed": elapsed, "total": total}}))
dist.destroy_process_group()
"""


def _torchrun(args, script_path, timeout=300):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29571", script_path] + args
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def test_two_ranks_gloo_partition_and_max_reduce(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    out = _torchrun([], str(script))
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    flat = sorted(x for part in res["gathered"] for x in part)
    assert flat == list(range(1000, 1006))
    assert set(res["gathered"][0]).isdisjoint(res["gathered"][1])
    assert res["elapsed"] == 2.0        # max over ranks, not the mean and not rank 0's own
    assert res["total"] == 6.0


def test_reference_arm_under_two_ranks_prints_one_line():
    """bench.py --impl reference under torchrun: rank 0 alone measures and prints; rank 1 exits 0 silently."""
    out = _torchrun(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                    os.path.join(ROOT, "bench.py"), timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["unit"] == "GH/s" and rec["value"] > 0
    assert rec["cpu_baseline"]["kind"] == "port" and rec["cpu_baseline"]["cores"] >= 1
    assert rec["e2e"]["h2d_bytes_per_step"] == 0 and rec["n_gpus"] == 2
