"""Drop-in for the reference ``miner.py``: same call surface, nonce loop on the GPU.

Kept from the reference (file:line are into /root/reference/miner.py):
  * ``calculate_merkle_root(transactions)``                         :15-18
  * ``mine_block(start_nonce=0, nonce_increment=1, mining_params)`` :30-114 -- wallet from
    ``sys.argv[1]``, rule and prefix derived with the reference's own expressions, the same
    progress prints, the same ``push_block`` POST and the final ``exit()``
  * ``mining_worker(start_nonce, nonce_increment, mining_params)``  :117-123
  * CLI ``python -m upow_b200.miner <wallet> [workers] [node_url]`` :126-156
Replaced: only the loop at :83-98, by one ``upow_search_min`` call into libupow_b200.so.
``workers`` now means "number of GPUs driven by this one process" (env UPOW_GPUS overrides).

There is no CPU fallback: without the CUDA library / a GPU this module raises.
"""
from __future__ import annotations

import os
import sys
import time
from datetime import datetime, timezone

from . import _lib
from .template import (ENDIAN, GENESIS_PREV_HASH, DifficultyRule, build_block_prefix, calculate_merkle_root,
                       string_to_bytes, with_timestamp)

__all__ = ["calculate_merkle_root", "mine_block", "mining_worker", "timestamp", "main"]

DEFAULT_NODE_URL = "http://localhost:3006/"
STALE_AFTER_S = 90      # miner.py:96-98
ROUND_LIMIT_S = 100     # miner.py:149-156

_context = None


def _node_url() -> str:
    """miner.py:21-27: argv[3] if given, else config.CORE_URL, else localhost."""
    if len(sys.argv) >= 4:
        return sys.argv[3].strip("/") + "/"
    try:
        import config  # the reference's config.py, if the operator has one on sys.path
        url = getattr(config, "CORE_URL", None)
        if url:
            return url
    except ImportError:
        pass
    return DEFAULT_NODE_URL


def timestamp() -> int:
    """UTC seconds, upow/helpers.py:37-38."""
    return int(datetime.now(timezone.utc).timestamp())


def get_context(n_gpus: int | None = None):
    """The process-wide GPU context (created on first use)."""
    global _context
    if _context is None:
        if n_gpus is None:
            n_gpus = int(os.environ.get("UPOW_GPUS", "1"))
        _context = _lib.Context(n_gpus)
    return _context


def set_context(ctx):
    global _context
    _context = ctx


def mine_block(start_nonce: int = 0, nonce_increment: int = 1, mining_params: dict = None):
    import requests  # only needed for the submit; imported late so tests can stub it

    miner_wallet = sys.argv[1]
    difficulty = mining_params["difficulty"]
    last_block = mining_params["last_block"]
    last_block["hash"] = last_block["hash"] if "hash" in last_block else GENESIS_PREV_HASH
    last_block["id"] = last_block["id"] if "id" in last_block else 0
    rule = DifficultyRule.derive(difficulty, last_block["hash"])

    wallet_bytes = string_to_bytes(miner_wallet)
    started = time.time()
    now = timestamp()
    tx_hashes = mining_params["pending_transactions_hashes"]
    merkle_root = calculate_merkle_root(tx_hashes)
    assert all(len(tx) == 64 for tx in tx_hashes)
    if start_nonce == 0:
        print(f"difficulty: {difficulty}")
        print(f'block number: {last_block["id"]}')
        print(f"Confirming {len(tx_hashes)} transactions")
    prefix = build_block_prefix(last_block["hash"], wallet_bytes, merkle_root, now, difficulty)

    ctx = get_context()
    ctx.set_template(0, prefix, rule.prev_hash, rule.n_int, rule.charset_count)
    max_rollovers = int(os.environ.get("UPOW_MAX_ROLLOVERS", "0"))
    rc, res = ctx.search_min(0, start_nonce, nonce_increment, time_limit_s=STALE_AFTER_S,
                             max_rollovers=max_rollovers)
    elapsed = max(time.time() - started, 1e-9)
    print(f"Worker {start_nonce + 1}: " + str(int(res.hashes_done / elapsed / 1000)) + "k hash/s")
    if rc == _lib.UPOW_EXHAUSTED and elapsed <= STALE_AFTER_S:
        # the reference dies here: nonce.to_bytes(4, ...) at 2**32 (miner.py:87)
        raise OverflowError("int too big to convert")
    if rc != _lib.UPOW_OK:
        return None  # stale or cancelled: nothing is submitted (miner.py:96-98)

    candidate = with_timestamp(prefix, res.timestamp) + res.nonce.to_bytes(4, ENDIAN)
    _submit(requests, candidate.hex(), tx_hashes, last_block["id"] + 1)
    exit()


def _submit(requests, block_content_hex: str, tx_hashes, block_no: int):
    """miner.py:99-113: print the candidate, POST it to <node>/push_block, print the reply."""
    print(block_content_hex)
    print(",".join(tx_hashes))
    response = requests.post(
        _node_url() + "push_block",
        json={"block_content": block_content_hex, "txs": tx_hashes, "block_no": block_no},
        timeout=20 + int((len(tx_hashes) or 1) / 3),
    )
    print(result := response.json())
    if result["ok"]:
        print("BLOCK MINED\n\n")
    return result


def mining_worker(start_nonce: int, nonce_increment: int, mining_params: dict):
    while True:
        mine_block(start_nonce, nonce_increment, mining_params)


def main():
    """One process drives all GPUs; otherwise the round structure of miner.py:126-156."""
    import requests

    n_gpus = int(os.environ.get("UPOW_GPUS", sys.argv[2] if len(sys.argv) >= 3 else 1))
    get_context(n_gpus)
    while True:
        print(f"Starting {n_gpus} workers")
        info = None
        while info is None:
            try:
                info = requests.get(_node_url() + "get_mining_info", timeout=5).json()["result"]
            except Exception as e:  # the reference retries forever, once a second
                print(e)
                time.sleep(1)
        # optional extras beyond the reference (both off by default so the CLI behaves like miner.py):
        #   UPOW_TEMPLATES=N     race N templates of this reply (timestamps x transaction subsets), multiplier.py
        #   UPOW_STALE_POLL_S=s  abandon the search as soon as the node's tip moves, stale.py
        n_templates = int(os.environ.get("UPOW_TEMPLATES", "1"))
        poll_s = float(os.environ.get("UPOW_STALE_POLL_S", "0"))
        watcher = None
        if poll_s > 0:
            from . import stale
            watcher = stale.StaleWatcher(get_context(), info["last_block"].get("id", 0),
                                         stale.http_tip_fetcher(_node_url()), poll_s)
            watcher.__enter__()
        try:
            if n_templates > 1:
                from . import multiplier
                out = multiplier.mine_any(get_context(), info, sys.argv[1], timestamp(), n_templates)
                if out is not None:
                    _submit(requests, out["block_content"], out["txs"], out["block_no"])
            else:
                mine_block(0, 1, info)
        except OverflowError:
            continue  # nonce space exhausted: fetch a fresh template (new timestamp)
        except SystemExit:
            continue  # the reference's worker exits after a submit and the parent starts the next round
        finally:
            if watcher is not None:
                watcher.__exit__(None, None, None)


if __name__ == "__main__":
    main()
