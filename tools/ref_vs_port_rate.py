#!/usr/bin/env python
"""Build-container only: time the reference's mine_block VERBATIM (imported from /root/reference through
oracle/ref_loader.py) and the oracle port `miner_loop` on the same core, on the same work -- the search for
the first valid nonce of template T0-v1 at a given difficulty.  Shows that the port bench.py times on the GPU
box (where the reference tree does not exist) costs what the reference costs."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader
from oracle import upow_oracle as O

d = float(sys.argv[1]) if len(sys.argv) > 1 else 6.3
out = []
for rep in range(3):
    t0 = time.perf_counter()
    r = ref_loader.run_mine_block(O.T0_ADDRESS_V1, O.t0_mining_params(d), O.T0_TIMESTAMP, 0, 1)
    t_ref = time.perf_counter() - t0
    nonce = int.from_bytes(bytes.fromhex(r["block_content"])[-4:], "little")
    prefix, rule = O.t0_prefix(d), O.Rule(d, O.T0_PREV_HASH)
    t0 = time.perf_counter()
    n2 = O.miner_loop(prefix, rule)
    t_port = time.perf_counter() - t0
    assert n2 == nonce
    out.append({"difficulty": d, "nonce": nonce, "reference_s": t_ref, "port_s": t_port,
                "reference_mhs": (nonce + 1) / t_ref / 1e6, "port_mhs": (nonce + 1) / t_port / 1e6,
                "port_over_reference_rate": t_ref / t_port})
    print(out[-1], flush=True)
json.dump(out, open(os.path.join(ROOT, "profiles", "r02_ref_vs_port.json"), "w"), indent=1)
