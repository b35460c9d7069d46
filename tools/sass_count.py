#!/usr/bin/env python
"""Count SASS instructions of the hot loop of each search kernel, per opcode and per pipe.

    python tools/sass_count.py upow_b200/libupow_b200.so [--kernel REGEX] [--json]

The hot loop of a kernel is taken to be the largest backward-branch region (the fully
unrolled per-nonce body).  Pipe classes follow the B300/B200 microarchitecture notes
(/opt/skills/guides/B300_MICROARCH.md, "Pipe rates"): IADD3/LOP3/SHF/PRMT/ISETP/SEL/... issue on
the alu pipe, IMAD*/FFMA/... on the fma pipe; everything else (branches, votes, uniform
datapath ops, loads) only costs an issue slot.
"""
import argparse
import collections
import json
import re
import subprocess
import sys

ALU = {"IADD3", "LOP3", "SHF", "PRMT", "ISETP", "SEL", "IABS", "LEA", "VIADD", "IMNMX", "VIMNMX", "PLOP3",
       "BMSK", "SGXT", "FLO", "POPC", "IADD", "VIADDMNMX", "MOV"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "IDP"}


def pipe_of(op):
    base = op.split(".")[0]
    if base in FMA:
        return "fma"
    if base in ALU:
        return "alu"
    if base.startswith("U") or base in ("R2UR", "S2UR"):
        return "uniform"
    return "other"


def parse(path):
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    funcs = {}
    cur = None
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", line)
        if m and cur is not None:
            addr = int(m.group(1), 16)
            ins = m.group(2).strip()
            funcs[cur].append((addr, ins))
    return funcs


def hot_loop(instrs):
    """(start, end) of the per-nonce loop: start from the largest backward-branch region and
    descend while it contains another loop at least half its size (the outer item loop wraps
    the inner nonce loop)."""
    loops = []
    for addr, ins in instrs:
        if re.match(r"(@!?U?P\d+\s+)?BRA", ins):
            t = re.search(r"0x([0-9a-f]+)", ins)
            if t:
                tgt = int(t.group(1), 16)
                if tgt <= addr:
                    loops.append((tgt, addr))
    if not loops:
        return None
    cur = max(loops, key=lambda l: l[1] - l[0])
    while True:
        inner = [l for l in loops if l != cur and cur[0] <= l[0] and l[1] <= cur[1]
                 and (l[1] - l[0]) * 2 >= (cur[1] - cur[0])]
        if not inner:
            return cur
        cur = max(inner, key=lambda l: l[1] - l[0])


def count(instrs, lo, hi):
    ops = collections.Counter()
    for addr, ins in instrs:
        if lo <= addr <= hi:
            parts = ins.split()
            op = parts[1] if parts[0].startswith("@") else parts[0]
            ops[op.split(".")[0] if not op.startswith("IMAD") else op.split(".")[0] + ("." + op.split(".")[1] if "." in op and op.split(".")[1] in ("HI", "WIDE", "SHL", "MOV", "IADD") else "")] += 1
    return ops


def analyse(path, kernel_re=None):
    out = {}
    for name, instrs in parse(path).items():
        if kernel_re and not re.search(kernel_re, name):
            continue
        loop = hot_loop(instrs)
        if not loop:
            continue
        ops = count(instrs, *loop)
        pipes = collections.Counter()
        for op, n in ops.items():
            pipes[pipe_of(op)] += n
        out[name] = {"loop": [hex(loop[0]), hex(loop[1])], "total": sum(ops.values()),
                     "pipes": dict(pipes), "ops": dict(ops.most_common()), "function_total": len(instrs)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("path")
    ap.add_argument("--kernel", default=None)
    ap.add_argument("--json", action="store_true")
    a = ap.parse_args()
    res = analyse(a.path, a.kernel)
    if a.json:
        json.dump(res, sys.stdout, indent=1)
        return
    for name, r in res.items():
        print(f"{name}\n  loop {r['loop']} total={r['total']} (function {r['function_total']}) pipes={r['pipes']}")
        print("  " + ", ".join(f"{k} {v}" for k, v in r["ops"].items()))


if __name__ == "__main__":
    main()
