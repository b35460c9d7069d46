#!/usr/bin/env python
"""Benchmark of the uPow nonce-search hot path (BASELINE.json metric: block hashes/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   (one rank per GPU)

One STEP = every GPU sweeps the full 2^32 nonce space of template T0-v1 at difficulty 5.0
(BASELINE configs[2], SURVEY 8(d) cfg 3: no early exit, exactly 2^32 hashes, all valid nonces
enumerated) for its own timestamp; timestamps advance by the cfg-4 rollover rule (ts+1 per
exhausted nonce space), so the ranks partition the (timestamp, nonce) space with no exchange:
weak scaling, no collective on the data path.  The first warm-up step of rank 0 runs timestamp
T0 itself and must reproduce SURVEY A.3 (4209 nonces, fingerprint 5ca83877...).

Prints ONE JSON line (rank 0).  `value` = hashes of all ranks / wall time of the K timed steps
(max over ranks, barrier + device sync on both sides); `e2e` = the same through the Python API
with host buffers (template bytes in, nonce list out); `roofline` = achieved vs. the integer
issue-rate roofline computed from the committed SASS count of the shipped kernel and opcode
rates measured live on this GPU; `cpu_baseline` = the reference's hashlib loop restated
(oracle port), one worker per host core, on a bounded sample.
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


# ------------------------------------------------------------------------------------------
# template T0 without touching the oracle package
# ------------------------------------------------------------------------------------------
def t0_prefix_and_rule(difficulty, timestamp):
    from upow_b200 import template as T
    h = lambda b: hashlib.sha256(b).hexdigest()
    prev = h(b"prev-block")
    wallet = h(b"addr-x") + h(b"addr-y")
    txs = [h(b"tx0"), h(b"tx1"), h(b"tx2")]
    prefix = T.build_block_prefix(prev, T.string_to_bytes(wallet), T.calculate_merkle_root(txs), timestamp,
                                  difficulty)
    return prefix, T.DifficultyRule.derive(difficulty, prev)


T0_TIMESTAMP = 1_700_000_000


# ------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.samples, self._stop, self._t = gpu_index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
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
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        reasons = []
        for idx, name in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap")):
            if any(s[idx].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        pw = [float(s[2]) for s in self.samples if s[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(self.samples[0][1]),
                "power_w_max": max(pw) if pw else None, "samples": len(self.samples), "reasons": reasons}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's hot loop restated (oracle port), one worker per core
# ------------------------------------------------------------------------------------------
def _cpu_worker(args):
    """Worker i of W walks nonces i, i+W, ... exactly like miner.py:139-148 + :83-89 (hashlib)."""
    start, inc, count, difficulty = args
    from oracle import upow_oracle as O
    prefix = O.t0_prefix(difficulty)
    rule = O.Rule(difficulty, O.T0_PREV_HASH)
    t0 = time.perf_counter()
    n = O.min_valid_nonce(prefix, rule, start, inc, start + inc * count)
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
        "config": {"workload": WORKLOAD_NAME, "note": "CPU arm: the reference's hashlib loop (oracle port of miner.py:83-89) "
                   "on the host cores; each step is a bounded sample of the workload"},
        "cpu_baseline": {"value": ghs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": ghs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


WORKLOAD_NAME = ("cfg3-sweep: template T0-v1 (138-byte message), difficulty 5.0, full 2^32 nonce sweep per GPU per step "
                 "enumerating all valid nonces; timestamp +1 per (step, rank) as in cfg 4's rollover rule")


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


def roofline(kernel_ghs, sm_mhz, sm_count, rates, sass):
    """H_peak = SMs * f / max(N_alu/R_alu, N_fma/R_fma, N_total/R_issue)  (SURVEY 8(d))."""
    if not sass or not rates or not sm_mhz or not sass.get("bench_kernel") or sass["bench_kernel"] not in sass.get("kernels", {}):
        return {"bound": "int_alu", "achieved": kernel_ghs, "peak": None, "unit": UNIT, "frac": None, "traffic": None}
    k = sass["kernels"][sass["bench_kernel"]]
    n_alu, n_fma, n_total = k["alu_per_hash"], k["fma_per_hash"], k["total_per_hash"]
    r_alu, r_fma, r_issue = rates["alu"], rates["fma"], rates["issue"]
    clk_per_warp_hash = max(n_alu / r_alu, n_fma / r_fma, n_total / r_issue)   # SM clocks per 32 hashes
    peak = sm_count * sm_mhz * 1e6 * 32.0 / clk_per_warp_hash / 1e9
    which = max((n_alu / r_alu, "alu pipe"), (n_fma / r_fma, "fma pipe"), (n_total / r_issue, "issue"))[1]
    return {"bound": "int_alu", "achieved": kernel_ghs, "peak": peak, "unit": UNIT, "frac": kernel_ghs / peak,
            "traffic": sass.get("dram_bytes_per_launch"),
            "binding": which, "sass_per_hash": {"alu": n_alu, "fma": n_fma, "total": n_total},
            "rates_warp_inst_per_clk_per_sm": {"alu": r_alu, "fma": r_fma, "issue": r_issue},
            "sm_mhz_in_kernel": sm_mhz, "sm_count": sm_count,
            "peak_source": "SASS count (profiles/sass_counts.json) x opcode rates measured live by upow_microbench",
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
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        except Exception:
            dist.init_process_group("gloo")

    import upow_b200
    if world > 1:
        devices = [local_rank]               # one rank per GPU
    else:
        devices = list(range(args.gpus))     # single process driving N GPUs through the library's own threads
    n_local = len(devices)
    ctx = upow_b200.Context(device_ids=devices)
    info = ctx.device_info(0)

    def barrier():
        if dist is not None:
            import torch
            if dist.get_backend() == "nccl":
                t = torch.zeros(1, device="cuda")
                dist.all_reduce(t)
                torch.cuda.synchronize()
            else:
                dist.barrier()

    from upow_b200 import sharding

    def allreduce_max(x: float) -> float:
        return sharding.reduce_elapsed(x, dist)

    def allreduce_sum(x: float) -> float:
        return sharding.reduce_sum(x, dist)

    n_total_gpus = world * n_local

    def step_timestamps(step_no):
        return sharding.step_timestamps(step_no, rank, world, n_local, T0_TIMESTAMP)

    prefix0, rule = t0_prefix_and_rule(SWEEP_DIFFICULTY, T0_TIMESTAMP)
    from upow_b200.template import with_timestamp

    def one_step_resident(step_no):
        """Templates already registered; per device one full sweep. Returns (hashes, kernel_ms list, counts)."""
        hashes, kms, counts, launches, mhz = 0, [], [], 0, []
        if n_local == 1:
            nonces, st = ctx.sweep(step_no % 60, 0, NONCE_SPACE)
            hashes += st.hashes_done; kms.append(st.kernel_ms); counts.append(len(nonces)); launches += st.launches
            mhz.append(st.sm_clock_mhz)
            return hashes, kms, counts, launches, mhz, nonces
        # single process, several GPUs: the library's tile scheduler hands one timestamp to each GPU
        rc, res = ctx.search_min(60 + step_no % 2, 0, 1, max_rollovers=n_local - 1, tile_nonces=NONCE_SPACE)
        assert rc == upow_b200.UPOW_EXHAUSTED
        return (res.hashes_done, [res.per_device_kernel_ms[i] for i in range(n_local)], [], res.launches,
                [res.sm_clock_mhz], [])

    total_steps = args.warmup + args.steps
    # ---- register the templates up front (inputs resident before the timed region) ----
    if n_local == 1:
        for s in range(min(total_steps, 60)):
            ts = step_timestamps(s)[0]
            ctx.set_template(s % 60, with_timestamp(prefix0, ts), rule.prev_hash, rule.n_int, rule.charset_count)
    else:
        # unreachable difficulty so that every GPU exhausts its timestamp (fixed work, SURVEY 8(d) cfg 4 variant)
        p16, r16 = t0_prefix_and_rule(16.0, T0_TIMESTAMP)
        for k in range(2):
            ctx.set_template(60 + k, p16, r16.prev_hash, r16.n_int, r16.charset_count)

    # ---- warm-up (first step of rank 0 is the SURVEY A.3 check) ----
    for s in range(args.warmup):
        out = one_step_resident(s)
        if s == 0 and rank == 0 and n_local == 1:
            nonces = out[5]
            fp = hashlib.sha256(b"".join(n.to_bytes(4, "little") for n in nonces)).hexdigest()
            if len(nonces) != A3_COUNT or fp != A3_FINGERPRINT:
                raise SystemExit(f"sweep parity FAILED: {len(nonces)} nonces, fingerprint {fp}")
        ctx.flush_l2()

    # ---- timed region: exactly K steps ----
    sampler = ClockSampler(devices[0])
    kernel_ms, hashes_local, launches_local, mhz_seen = [], 0, 0, []
    barrier()
    with sampler:
        t0 = time.perf_counter()
        for s in range(args.warmup, total_steps):
            h, kms, _, ln, mhz, _ = one_step_resident(s)
            hashes_local += h; kernel_ms += kms; launches_local += ln; mhz_seen += [m for m in mhz if m]
            ctx.flush_l2()   # cold L2 for the next step (the kernel reads no HBM inputs)
        elapsed_local = time.perf_counter() - t0
    barrier()
    elapsed = allreduce_max(elapsed_local)
    hashes_all = allreduce_sum(float(hashes_local))
    launches_all = int(allreduce_sum(float(launches_local)))
    value = hashes_all / elapsed / 1e9

    # ---- e2e: host bytes in (template), host list out (nonces), every step ----
    barrier()
    t0 = time.perf_counter()
    e2e_hashes, d2h, h2d = 0, 0, 0
    for s in range(args.steps):
        if n_local == 1:
            ts = step_timestamps(total_steps + s)[0]
            ctx.set_template(0, with_timestamp(prefix0, ts), rule.prev_hash, rule.n_int, rule.charset_count)
            nonces, st = ctx.sweep(0, 0, NONCE_SPACE)
            checksum = sum(nonces) & 0xFFFFFFFF  # the step's result is read on the host
            e2e_hashes += st.hashes_done
            h2d += len(prefix0) + 32 + 8 + 48       # template bytes + rule + DevResult init
            d2h += 4 * len(nonces) + 64             # nonce list + DevResult
        else:
            p16, r16 = t0_prefix_and_rule(16.0, T0_TIMESTAMP + 1000 + s * n_local)
            ctx.set_template(62, p16, r16.prev_hash, r16.n_int, r16.charset_count)
            rc, res = ctx.search_min(62, 0, 1, max_rollovers=n_local - 1, tile_nonces=NONCE_SPACE)
            e2e_hashes += res.hashes_done
            h2d += (len(p16) + 32 + 8 + 48) * n_local
            d2h += 64 * n_local
    e2e_local = time.perf_counter() - t0
    barrier()
    e2e_elapsed = allreduce_max(e2e_local)
    e2e_value = allreduce_sum(float(e2e_hashes)) / e2e_elapsed / 1e9

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (search_kernel<V1, ENUM>) on this rank's GPU ----
    kernel_ghs = (NONCE_SPACE / (statistics.mean(kernel_ms) / 1e3)) / 1e9 if kernel_ms else None
    sm_mhz = statistics.median(mhz_seen) if mhz_seen else None
    rates, mb = (None, None)
    try:
        rates, mb = measure_rates(devices[0])
    except Exception as e:  # microbench failure must not kill the bench line
        print(f"microbench failed: {e}", file=sys.stderr)
    roof = roofline(kernel_ghs, sm_mhz, info.sm_count, rates, load_sass_counts())
    roof["kernel_ms_avg"] = statistics.mean(kernel_ms) if kernel_ms else None

    cpu = None
    if world * n_local == 1 and not args.no_cpu_baseline:
        r = cpu_reference_rate(1_500_000)   # ~2 s per worker
        cpu = {"value": r["rate_hs"] / 1e9, "unit": UNIT, "cores": r["cores"], "kind": "port",
               "sample": f"{r['hashes']} nonces of the same template (unreachable difficulty), hashlib loop restating "
                         f"miner.py:83-89, one process per core striding like miner.py:139-148, {r['max_worker_s']:.1f} s"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_total_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": elapsed / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": WORKLOAD_NAME, "hashes_per_gpu_per_step": NONCE_SPACE,
                   "parallelism": f"{n_total_gpus} independent nonce-space shards, no collective",
                   "launcher": "torchrun, one rank per GPU" if world > 1 else "single process",
                   "l2": "flushed between steps (256 MiB memset inside the timed region); the kernel reads no HBM inputs",
                   "timing": "value: host clock around the K steps, every step ends in a stream synchronise, barrier on both "
                             "sides, max over ranks; roofline.achieved: CUDA events on the library's stream around each kernel",
                   "device": info.name.decode(), "grid_blocks": info.search_grid_blocks},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // max(args.steps, 1),
                "d2h_bytes_per_step": d2h // max(args.steps, 1)},
        "gpu_launches": launches_all,
        "roofline": roof,
        "cpu_baseline": cpu,
        "clocks": sampler.summary(),
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
