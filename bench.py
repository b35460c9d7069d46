#!/usr/bin/env python
"""Benchmark of the uPow nonce-search hot path (BASELINE.json metric: block hashes/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   (one rank per GPU)

N = 1.  One STEP = one full 2^32 nonce sweep of template T0-v1 at difficulty 5.0 (BASELINE configs[2],
SURVEY 8(d) cfg 3: no early exit, exactly 2^32 hashes, all valid nonces enumerated), a new timestamp every step.
The first warm-up step runs timestamp T0 itself and must reproduce SURVEY A.3 (4209 nonces, 5ca83877...).

N > 1 (torchrun, one rank per GPU).  Two things are measured, in this order:
  (i)  every rank sweeps its own timestamps on its own GPU, exactly like N = 1 (N processes that share
       nothing) -> config.independent_ranks_ghs;
  (ii) the other ranks park in a CPU (gloo) barrier and rank 0 alone drives ALL N GPUs through ONE library
       context -- the path north_star names: one nonce space shared by the GPUs (lgroup interleave), found
       words travelling between the GPUs, exact minimum, timestamp rollover.  One STEP here = 4 N full
       timestamps at the unreachable difficulty 16.0 (fixed work, 4 N * 2^32 hashes, SURVEY 8(d) cfg 4
       "fixed-work variant").  THIS is `value` for N > 1.
After the timed steps rank 0 also runs, on the N GPUs it drives (N = 1 included): BASELINE config 4 for real
(T0-v1 at 9.5 with rollover; the answer is asserted), config 5 (4096 templates at 6.0; the per-template
minima are asserted against the CPU-oracle fixture's fingerprint) and config 2 / config 1 as time-to-solution.

Prints ONE JSON line (rank 0).  `e2e` = the `value` loop again, with the template bytes handed over from host
memory and the result read back on the host every step; `roofline` = achieved vs. the integer issue-rate
roofline of the shipped kernel's SASS and of the algorithmic floor; `cpu_baseline` = the reference's hashlib
loop restated (oracle port), one worker per host core, on a bounded sample.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NONCE_SPACE = 1 << 32
SWEEP_DIFFICULTY = 5.0
A3_COUNT = 4209
A3_FINGERPRINT = "5ca83877d72644292a359a6a2864d9411523a9c01a2f2a0ec6aea1ee3c9d742f"
METRIC = "block hashes/sec"
UNIT = "GH/s"
T0_TIMESTAMP = 1_700_000_000

# BASELINE config 4 at difficulty 9.5: first valid (timestamp, nonce) in search order, and its digest.
# Found identically by the plain-CUDA kernel, the tuned kernel and every GPU count (profiles/r01_configs_8gpu.json);
# re-hashed with hashlib below.  (The 9.0 answer is pinned by the CPU oracle: tests/golden/cfg4_t0v1_d9.json.)
CFG4_DIFFICULTY = 9.5
CFG4_ANSWER = (1_700_000_052, 4_084_426_313)
CFG4_DIGEST_PREFIX = "9053745dc"
# BASELINE config 5: sha256 of the 4096 per-template minimum nonces (u32 LE, template order), from the CPU oracle
# (tests/golden/cfg5_min_nonces.json, written by oracle/gen_golden.py --cfg5).
CFG5_DIFFICULTY = 6.0
CFG5_TEMPLATES = 4096

# Algorithmic floor of alu-pipe instructions per hash (DESIGN.md "roofline anchors"): rotates/shifts and
# three-input logic ops that no pipe but the alu pipe can execute, after all template-specific folding.
ALU_FLOOR_PER_HASH = {1: 922, 2: 818}


def _h(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()


# ------------------------------------------------------------------------------------------
# templates without touching the oracle package
# ------------------------------------------------------------------------------------------
def t0_prefix_and_rule(difficulty, timestamp):
    from upow_b200 import template as T
    prev = _h(b"prev-block")
    wallet = _h(b"addr-x") + _h(b"addr-y")
    txs = [_h(b"tx0"), _h(b"tx1"), _h(b"tx2")]
    prefix = T.build_block_prefix(prev, T.string_to_bytes(wallet), T.calculate_merkle_root(txs), timestamp,
                                  difficulty)
    return prefix, T.DifficultyRule.derive(difficulty, prev)


def cfg5_templates(difficulty=CFG5_DIFFICULTY, n=CFG5_TEMPLATES):
    """SURVEY 8(d) cfg 5: template i varies the address and the transaction list."""
    from upow_b200 import template as T
    prev = _h(b"prev-block")
    rule = T.DifficultyRule.derive(difficulty, prev)
    out = []
    for i in range(n):
        le = i.to_bytes(4, "little")
        wallet = _h(b"addr-x" + le) + _h(b"addr-y" + le)
        txs = [_h(b"tx" + le + bytes([j])) for j in range(3)]
        prefix = T.build_block_prefix(prev, T.string_to_bytes(wallet), T.calculate_merkle_root(txs), T0_TIMESTAMP, difficulty)
        out.append((prefix, rule.prev_hash, rule.n_int, rule.charset_count))
    return out


def cfg5_expected_fingerprint():
    try:
        with open(os.path.join(ROOT, "tests", "golden", "cfg5_min_nonces.json")) as f:
            return json.load(f)["sha256_u32le_in_template_order"]
    except (OSError, KeyError, ValueError):
        return None


# ------------------------------------------------------------------------------------------
# reductions over ranks (torch.distributed is plumbing: barrier, max of the elapsed time, sum of hashes)
# ------------------------------------------------------------------------------------------
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


# ------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_indices):
        self.gpus, self.samples, self._stop, self._t = list(gpu_indices), [], threading.Event(), None

    def _run(self):
        ids = ",".join(str(g) for g in self.gpus)
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={ids}", f"--query-gpu={self.FIELDS}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                for line in out.strip().splitlines():
                    parts = [p.strip() for p in line.split(",")]
                    if len(parts) >= 8:
                        self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        num = lambda s: float(s) if s.replace(".", "").isdigit() else None
        sm = [num(s[1]) for s in self.samples if num(s[1]) is not None]
        reasons = []
        for idx, name in ((4, "hw_slowdown"), (5, "hw_thermal_slowdown"), (6, "sw_thermal_slowdown"), (7, "sw_power_cap")):
            if any(s[idx].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        pw = [num(s[3]) for s in self.samples if num(s[3]) is not None]
        per_gpu = {}
        for g in self.gpus:
            v = [num(s[1]) for s in self.samples if s[0] == str(g) and num(s[1]) is not None]
            if v:
                per_gpu[str(g)] = statistics.median(v)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": num(self.samples[0][2]),
                "power_w_max": max(pw) if pw else None, "samples": len(self.samples), "reasons": reasons,
                "sm_mhz_per_gpu": per_gpu, "sm_mhz_min_gpu": min(per_gpu.values()) if per_gpu else None}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's hot loop restated (oracle port), one worker per core
# ------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """Worker i of W walks nonces i, i+W, ... exactly like miner.py:139-148 + :83-98 (hashlib, progress test)."""
    start, inc, count, difficulty = args
    from oracle import upow_oracle as O
    prefix = O.t0_prefix(difficulty)
    rule = O.Rule(difficulty, O.T0_PREV_HASH)
    t0 = time.perf_counter()
    n = O.miner_loop(prefix, rule, start, inc, start + inc * count)
    dt = time.perf_counter() - t0
    assert n is None  # difficulty 16.0 is unreachable: every worker hashes exactly `count` nonces
    return dt


def cpu_reference_rate(per_worker: int, workers: int | None = None):
    workers = workers or os.cpu_count() or 1
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(workers) as pool:
        dts = pool.map(_cpu_worker, [(i, workers, per_worker, 16.0) for i in range(workers)])
    wall = time.perf_counter() - t0
    total = per_worker * workers
    return {"hashes": total, "wall_s": wall, "max_worker_s": max(dts), "rate_hs": total / max(dts), "cores": workers}


CPU_ARM_NOTE = ("CPU arm: the reference's loop miner.py:83-98 restated statement by statement over hashlib "
                "(oracle port, incl. the per-iteration progress test); the reference is pure Python and its tree "
                "does not exist on the GPU box, so the port is what can be timed there")


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_worker = 1_000_000  # ~0.7 s of hashlib per worker per step
    for _ in range(args.warmup):
        cpu_reference_rate(per_worker // 5, cores)
    t0 = time.perf_counter()
    hashes = 0
    for _ in range(args.steps):
        r = cpu_reference_rate(per_worker, cores)
        hashes += r["hashes"]
    wall = time.perf_counter() - t0
    ghs = hashes / wall / 1e9
    sample = f"{per_worker} nonces per worker per step, {cores} workers striding like miner.py:139-148, unreachable difficulty 16.0"
    line = {
        "impl": "reference", "metric": METRIC, "value": ghs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": wall / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": workload_name(args.gpus), "note": CPU_ARM_NOTE + "; each step is a bounded sample of the workload"},
        "cpu_baseline": {"value": ghs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": ghs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_name(n_gpus):
    if n_gpus <= 1:
        return ("cfg3-sweep: template T0-v1 (138-byte message), difficulty 5.0, full 2^32 nonce sweep per step "
                "enumerating all valid nonces; timestamp +1 per step")
    return (f"cfg4-fixed-work: template T0-v1 at the unreachable difficulty 16.0, {4 * n_gpus} timestamps x 2^32 nonces per step "
            f"(rollover ts+1 on exhaustion), ONE library context driving {n_gpus} GPUs that share every nonce space")


# ------------------------------------------------------------------------------------------
# roofline from the committed SASS count + live opcode rates
# ------------------------------------------------------------------------------------------
def load_sass_counts():
    path = os.path.join(ROOT, "profiles", "sass_counts.json")
    try:
        with open(path) as f:
            return json.load(f)
    except OSError:
        return None


def roofline(kernel_ghs, sm_mhz, sm_count, rates, sass, layout=1):
    """H_peak = SMs * f * 32 / max(N_alu/R_alu, N_fma/R_fma, N_total/R_issue)  (SURVEY 8(d)).

    Two anchors beside it, so that the fraction cannot be flattered by a kernel that merely carries more
    alu instructions: frac_vs_floor uses the algorithmic alu-pipe floor instead of the shipped count, and
    frac_issue is the share of all issue slots the shipped instruction stream could fill."""
    if not sass or not rates or not sm_mhz or not sass.get("bench_kernel") or sass["bench_kernel"] not in sass.get("kernels", {}):
        return {"bound": "int_alu", "achieved": kernel_ghs, "peak": None, "unit": UNIT, "frac": None, "traffic": None}
    k = sass["kernels"][sass["bench_kernel"]]
    n_alu, n_fma, n_total = k["alu_per_hash"], k["fma_per_hash"], k["total_per_hash"]
    r_alu, r_fma, r_issue = rates["alu"], rates["fma"], rates["issue"]
    hz = sm_count * sm_mhz * 1e6 * 32.0 / 1e9            # lane-instructions per ns, all SMs
    clk_per_warp_hash = max(n_alu / r_alu, n_fma / r_fma, n_total / r_issue)   # SM clocks per 32 hashes
    peak = hz / clk_per_warp_hash
    which = max((n_alu / r_alu, "alu pipe"), (n_fma / r_fma, "fma pipe"), (n_total / r_issue, "issue"))[1]
    floor = ALU_FLOOR_PER_HASH[layout]
    peak_floor = hz * r_alu / floor
    peak_issue = hz * r_issue / n_total
    return {"bound": "int_alu", "achieved": kernel_ghs, "peak": peak, "unit": UNIT, "frac": kernel_ghs / peak,
            "traffic": sass.get("dram_bytes_per_launch"),
            "binding": which, "sass_per_hash": {"alu": n_alu, "fma": n_fma, "total": n_total},
            "alu_floor_per_hash": floor, "alu_over_floor": n_alu / floor,
            "peak_at_alu_floor": peak_floor, "frac_vs_floor": kernel_ghs / peak_floor,
            "peak_issue": peak_issue, "frac_issue": kernel_ghs / peak_issue,
            "rates_warp_inst_per_clk_per_sm": {"alu": r_alu, "fma": r_fma, "issue": r_issue},
            "sm_mhz_in_kernel": sm_mhz, "sm_count": sm_count,
            "peak_source": "SASS count (profiles/sass_counts.json) x opcode rates measured live by upow_microbench; "
                           "peak_at_alu_floor uses the algorithmic floor of alu-only instructions (DESIGN.md)",
            "algorithmic_bytes_per_hash": 0}


def measure_rates(device):
    import upow_b200
    mb = {r["name"]: r for r in upow_b200.microbench(device, iters=1500)}
    # alu pipe: LOP3 / SHF / IADD3 with three live register operands; fma pipe: IMAD; issue: the
    # best interleaved alu+fma stream (SHF + IMAD)
    alu = max(mb[n]["warp_inst_per_clk_per_sm"] for n in ("LOP3", "SHF.rot", "LOP3 3 reg reads", "IADD3 3 reg reads"))
    fma = max(mb[n]["warp_inst_per_clk_per_sm"] for n in ("IMAD", "IMAD 2 reg reads (imm)"))
    issue = max(mb[n]["warp_inst_per_clk_per_sm"] for n in ("mix SHF+IMAD 1:1", "mix SHF(1r)+IMAD(3r)"))
    return {"alu": alu, "fma": fma, "issue": issue}, mb


# ------------------------------------------------------------------------------------------
# the BASELINE configs run through the library's own multi-GPU path (rank 0, all its GPUs)
# ------------------------------------------------------------------------------------------
def run_cfg4(ctx, n_gpus):
    import upow_b200
    prefix, rule = t0_prefix_and_rule(CFG4_DIFFICULTY, T0_TIMESTAMP)
    ctx.set_template(40, prefix, rule.prev_hash, rule.n_int, rule.charset_count)
    t0 = time.perf_counter()
    rc, res = ctx.search_min(40, max_rollovers=4096)
    wall = time.perf_counter() - t0
    got = (res.timestamp, res.nonce)
    from upow_b200.template import with_timestamp
    block = with_timestamp(prefix, res.timestamp) + res.nonce.to_bytes(4, "little")
    ok = (rc == upow_b200.UPOW_OK and got == CFG4_ANSWER and hashlib.sha256(block).hexdigest() == res.digest_hex
          and res.digest_hex.startswith(CFG4_DIGEST_PREFIX))
    if not ok:
        raise SystemExit(f"cfg4 parity FAILED on {n_gpus} GPUs: rc={rc} answer={got} digest={res.digest_hex}")
    return {"difficulty": CFG4_DIFFICULTY, "ghs": res.hashes_done / wall / 1e9, "wall_s": wall, "hashes": res.hashes_done,
            "answer": {"timestamp": res.timestamp, "nonce": res.nonce, "digest": res.digest_hex}, "rollovers": res.rollovers,
            "overshoot": res.hashes_done / (res.rollovers * NONCE_SPACE + res.nonce + 1), "launches": res.launches,
            "answer_asserted": True}


def run_cfg5(ctx, n_gpus, templates):
    arr = ctx.template_array(templates)      # the caller's host buffer, in the C layout of include/upow_b200.h
    items, stats = ctx.search_batch(arr)
    # the library call itself, host buffers in, host buffers out: template preparation, uploads, search, download,
    # host cross-check of every result (Python's own list building is not part of the path)
    wall = stats.elapsed_s
    if not all(found for found, _, _ in items):
        raise SystemExit("cfg5: a template has no valid nonce")
    fp = hashlib.sha256(b"".join(n.to_bytes(4, "little") for _, n, _ in items)).hexdigest()
    want = cfg5_expected_fingerprint()
    if want is not None and fp != want:
        raise SystemExit(f"cfg5 parity FAILED on {n_gpus} GPUs: fingerprint {fp} != CPU-oracle fixture {want}")
    needed = sum(n + 1 for _, n, _ in items)
    return {"difficulty": CFG5_DIFFICULTY, "templates": len(templates), "ghs": stats.hashes_done / wall / 1e9,
            "ghs_kernel": stats.hashes_done / (stats.kernel_ms / 1e3) / 1e9 if stats.kernel_ms else None,
            "wall_s": wall, "kernel_ms": stats.kernel_ms, "hashes": stats.hashes_done, "overshoot": stats.hashes_done / needed,
            "fingerprint": fp, "fingerprint_asserted": want is not None, "launches": stats.launches}


def run_short(ctx, difficulty, reps=20):
    """Time to solution of an easy search (BASELINE configs[1] = 6.3, configs[0] = 4.0) through Context.search_min."""
    prefix, rule = t0_prefix_and_rule(difficulty, T0_TIMESTAMP)
    ctx.set_template(41, prefix, rule.prev_hash, rule.n_int, rule.charset_count)
    ctx.search_min(41)
    el, km, hs, nonce = [], [], [], None
    for _ in range(reps):
        t0 = time.perf_counter()
        rc, res = ctx.search_min(41)
        el.append((time.perf_counter() - t0) * 1e3)
        km.append(res.kernel_ms)
        hs.append(res.hashes_done)
        nonce = res.nonce
        assert rc == 0
    return {"difficulty": difficulty, "nonce": nonce, "elapsed_ms": statistics.median(el), "kernel_ms": statistics.median(km),
            "overhead_ms": statistics.median(el) - statistics.median(km), "hashes_done": int(statistics.median(hs)),
            "overshoot": statistics.median(hs) / (nonce + 1), "reps": reps}


# ------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the cfg 2/4/5 runs after the timed steps")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    dist, gloo = None, None
    if world > 1:
        import datetime
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        except Exception:
            dist.init_process_group("gloo")
        # a CPU-side group for parking: ranks waiting in an NCCL collective would spin kernels on their GPUs
        gloo = dist.new_group(backend="gloo", timeout=datetime.timedelta(minutes=30))

    import upow_b200
    from upow_b200 import sharding
    from upow_b200.template import with_timestamp

    def barrier():
        if dist is not None:
            import torch
            if dist.get_backend() == "nccl":
                t = torch.zeros(1, device="cuda")
                dist.all_reduce(t)
                torch.cuda.synchronize()
            else:
                dist.barrier()

    def park():
        if dist is not None:
            dist.barrier(group=gloo)

    # ================= phase (i): one GPU per process, cfg-3 sweeps =================
    my_dev = local_rank if world > 1 else 0
    ctx = upow_b200.Context(device_ids=[my_dev])
    info = ctx.device_info(0)
    prefix0, rule = t0_prefix_and_rule(SWEEP_DIFFICULTY, T0_TIMESTAMP)
    total_steps = args.warmup + 2 * args.steps

    def step_ts(step_no):
        return sharding.step_timestamps(step_no, rank, world, 1, T0_TIMESTAMP)[0]

    for s in range(min(total_steps, 60)):   # templates registered up front: inputs resident before the timed region
        ctx.set_template(s % 60, with_timestamp(prefix0, step_ts(s)), rule.prev_hash, rule.n_int, rule.charset_count)

    def sweep_step(slot):
        nonces, st = ctx.sweep(slot, 0, NONCE_SPACE)
        return nonces, st

    for s in range(args.warmup):
        nonces, _ = sweep_step(s % 60)
        if s == 0 and rank == 0:
            fp = hashlib.sha256(b"".join(n.to_bytes(4, "little") for n in nonces)).hexdigest()
            if len(nonces) != A3_COUNT or fp != A3_FINGERPRINT:
                raise SystemExit(f"sweep parity FAILED: {len(nonces)} nonces, fingerprint {fp}")
        ctx.flush_l2()

    sampler1 = ClockSampler([my_dev])
    kernel_ms, hashes_local, launches_local, mhz_seen = [], 0, 0, []
    barrier()
    with sampler1:
        t0 = time.perf_counter()
        for s in range(args.warmup, args.warmup + args.steps):
            _, st = sweep_step(s % 60)
            hashes_local += st.hashes_done; kernel_ms.append(st.kernel_ms); launches_local += st.launches
            if st.sm_clock_mhz:
                mhz_seen.append(st.sm_clock_mhz)
            ctx.flush_l2()   # cold L2 for the next step (the kernel reads no HBM inputs)
        elapsed_local = time.perf_counter() - t0
    barrier()
    sweep_elapsed = reduce_elapsed(elapsed_local, dist)
    sweep_hashes = reduce_sum(float(hashes_local), dist)
    sweep_launches = int(reduce_sum(float(launches_local), dist))
    sweep_value = sweep_hashes / sweep_elapsed / 1e9

    # e2e of the same loop: template bytes from host memory every step, nonce list read on the host, same flush
    barrier()
    t0 = time.perf_counter()
    e2e_hashes, d2h, h2d = 0, 0, 0
    for s in range(args.steps):
        ts = step_ts(args.warmup + args.steps + s)
        ctx.set_template(0, with_timestamp(prefix0, ts), rule.prev_hash, rule.n_int, rule.charset_count)
        nonces, st = sweep_step(0)
        checksum = sum(nonces) & 0xFFFFFFFF  # the step's result is read on the host
        e2e_hashes += st.hashes_done
        h2d += len(prefix0) + 32 + 8 + 128       # template bytes + rule + launch state
        d2h += 4 * len(nonces) + 128             # nonce list + launch state
        ctx.flush_l2()
    e2e_local = time.perf_counter() - t0
    barrier()
    sweep_e2e = reduce_sum(float(e2e_hashes), dist) / reduce_elapsed(e2e_local, dist) / 1e9
    sweep_h2d, sweep_d2h = h2d // max(args.steps, 1), d2h // max(args.steps, 1)

    if rank != 0:
        ctx.close()
        park()      # rank 0 now drives every GPU alone; wait on the CPU until it is done
        dist.destroy_process_group()
        return

    # ================= phase (ii): rank 0 drives all N GPUs through one context =================
    n_gpus = world if world > 1 else max(args.gpus, 1)
    multi = None
    if n_gpus > 1:
        p16, r16 = t0_prefix_and_rule(16.0, T0_TIMESTAMP)
        # the same fixed-work search on ONE GPU (4 timestamps), for a like-for-like single-GPU figure: the search
        # kernel (MODE_MIN, with its limit polls) is 0.5 % slower than the enumerating sweep kernel timed at N = 1
        ctx.set_template(49, p16, r16.prev_hash, r16.n_int, r16.charset_count)
        ctx.search_min(49, max_rollovers=1)
        t0 = time.perf_counter()
        rc, res1 = ctx.search_min(49, max_rollovers=3)
        fixed_1gpu = res1.hashes_done / (time.perf_counter() - t0) / 1e9
        ctx.close()
        ctx = upow_b200.Context(device_ids=list(range(n_gpus)))
        per_step_ts = 4 * n_gpus
        for k in range(4):
            ctx.set_template(50 + k, with_timestamp(p16, T0_TIMESTAMP + 1000 * k), r16.prev_hash, r16.n_int, r16.charset_count)

        def fixed_step(slot):
            rc, res = ctx.search_min(slot, max_rollovers=per_step_ts - 1)
            if rc != upow_b200.UPOW_EXHAUSTED or res.hashes_done != per_step_ts * NONCE_SPACE:
                raise SystemExit(f"fixed-work step: rc={rc} hashes={res.hashes_done}")
            return res

        for s in range(args.warmup):
            fixed_step(50 + s % 4)
            ctx.flush_l2()
        sampler = ClockSampler(range(n_gpus))
        m_hashes, m_launches, m_dev_ms = 0, 0, []
        with sampler:
            t0 = time.perf_counter()
            for s in range(args.steps):
                res = fixed_step(50 + s % 4)
                m_hashes += res.hashes_done; m_launches += res.launches
                m_dev_ms.append(max(res.per_device_kernel_ms[i] for i in range(n_gpus)))
                ctx.flush_l2()
            m_elapsed = time.perf_counter() - t0
        # e2e: the template comes from host bytes every step, the result struct is read on the host
        t0 = time.perf_counter()
        me_hashes = 0
        for s in range(args.steps):
            ctx.set_template(54, with_timestamp(p16, T0_TIMESTAMP + 5000 + 100 * s), r16.prev_hash, r16.n_int, r16.charset_count)
            res = fixed_step(54)
            me_hashes += res.hashes_done
            ctx.flush_l2()
        me_elapsed = time.perf_counter() - t0
        multi = {"value": m_hashes / m_elapsed / 1e9, "elapsed": m_elapsed, "launches": m_launches,
                 "e2e": me_hashes / me_elapsed / 1e9, "device_ms_per_step": statistics.mean(m_dev_ms),
                 "h2d": (len(p16) + 32 + 8) + 128 * per_step_ts * n_gpus, "d2h": 128 * per_step_ts * n_gpus,
                 "clocks": sampler.summary(), "fixed_1gpu": fixed_1gpu,
                 "peer_devices": [ctx.device_info(i).peer_devices for i in range(n_gpus)]}

    # ---- the BASELINE configs through the same context (answers asserted) ----
    extra = {}
    if not args.no_configs:
        extra["cfg4"] = run_cfg4(ctx, n_gpus)
        extra["cfg5"] = run_cfg5(ctx, n_gpus, cfg5_templates())
        if n_gpus == 1:
            extra["cfg2"] = run_short(ctx, 6.3)
            extra["cfg1"] = run_short(ctx, 4.0)

    # ---- roofline of the dominant kernel (search_kernel_tuned<V1, ENUM>) on this rank's GPU ----
    kernel_ghs = (NONCE_SPACE / (statistics.mean(kernel_ms) / 1e3)) / 1e9 if kernel_ms else None
    sm_mhz = statistics.median(mhz_seen) if mhz_seen else None
    rates, mb = (None, None)
    try:
        rates, mb = measure_rates(0)
    except Exception as e:  # microbench failure must not kill the bench line
        print(f"microbench failed: {e}", file=sys.stderr)
    roof = roofline(kernel_ghs, sm_mhz, info.sm_count, rates, load_sass_counts())
    roof["kernel_ms_avg"] = statistics.mean(kernel_ms) if kernel_ms else None
    ctx.close()
    if dist is not None:
        park()      # release the parked ranks

    cpu = None
    if n_gpus == 1 and world == 1 and not args.no_cpu_baseline:
        r = cpu_reference_rate(1_500_000)   # ~2 s per worker
        cpu = {"value": r["rate_hs"] / 1e9, "unit": UNIT, "cores": r["cores"], "kind": "port",
               "sample": f"{r['hashes']} nonces of the same template (unreachable difficulty), hashlib loop restating "
                         f"miner.py:83-98, one process per core striding like miner.py:139-148, {r['max_worker_s']:.1f} s"}

    if multi is None:
        value, elapsed, launches, e2e_value = sweep_value, sweep_elapsed, sweep_launches, sweep_e2e
        h2d_b, d2h_b, clocks = sweep_h2d, sweep_d2h, sampler1.summary()
        parallelism = "1 GPU"
    else:
        value, elapsed, launches, e2e_value = multi["value"], multi["elapsed"], multi["launches"], multi["e2e"]
        h2d_b, d2h_b, clocks = multi["h2d"], multi["d2h"], multi["clocks"]
        parallelism = (f"one context, {n_gpus} GPUs share every nonce space (lgroup interleave), finds travel GPU to GPU, "
                       "no collective")
    config = {"workload": workload_name(n_gpus), "hashes_per_gpu_per_step": NONCE_SPACE if multi is None else 4 * NONCE_SPACE,
              "parallelism": parallelism,
              "launcher": "torchrun, one rank per GPU; ranks 1.. parked in a gloo barrier while rank 0 drives all GPUs" if world > 1 else "single process",
              "l2": "flushed between steps (256 MiB memset inside the timed region); the kernel reads no HBM inputs",
              "timing": "value: host clock around the K steps, every step ends with the library call returning (all streams "
                        "drained), barrier on both sides; roofline.achieved: CUDA events on the library's stream around each kernel",
              "device": info.name.decode(), "grid_blocks": info.search_grid_blocks,
              "independent_ranks_ghs": sweep_value, "independent_ranks_e2e_ghs": sweep_e2e}
    if multi is not None:
        config["device_ms_per_step"] = multi["device_ms_per_step"]
        config["peer_devices"] = multi["peer_devices"]
        config["fixed_work_1gpu_ghs"] = multi["fixed_1gpu"]
        config["scaling_vs_same_search_on_1gpu"] = multi["value"] / (n_gpus * multi["fixed_1gpu"])
    config.update(extra)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": config,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_b, "d2h_bytes_per_step": d2h_b},
        "gpu_launches": launches,
        "roofline": roof,
        "cpu_baseline": cpu,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if mb is not None:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "microbench_last.json"), "w") as f:
            json.dump(mb, f, indent=1)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
