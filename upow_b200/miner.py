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

Where a B200 forces a difference: the reference never exhausts a nonce space (2^32 hashes take it more than an
hour), a B200 does so in 0.23 s.  The reference's answer to exhaustion is to die with OverflowError (:87) and be
restarted with a fresh ``timestamp()`` (:149-156) -- which, inside the same wall-clock second, is the SAME
template.  So here the search rolls the block timestamp BACKWARDS (now, now-1, ... down to
last_block.timestamp+1: every one of those is acceptable to the node, upow/manager.py:445-460, none lies in
the future) inside one library call, the spaces already searched for a chain tip are remembered and never
searched again, and ``get_mining_info`` (rate limited to 30/minute, upow/node/main.py:675-676) is asked at most
once every two seconds.

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
MINING_INFO_MIN_PERIOD_S = 2.0   # 30 requests/minute at the node (upow/node/main.py:675-676)
PROGRESS_PERIOD_S = 5.0          # a reference worker prints its rate every 5e6 hashes ~ 5 s (miner.py:89-95)

_context = None
# (prev_hash, wallet, merkle_root, difficulty, start, stride) -> (newest, oldest) timestamp whose nonce space
# has been searched to the end without a find
_searched: dict = {}


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

    # which timestamps are still worth searching for this template: now, now-1, ... down to the block after the
    # previous one, minus what earlier calls already searched to the end
    key = (last_block["hash"], miner_wallet, merkle_root, str(difficulty), start_nonce, nonce_increment)
    last_ts = last_block.get("timestamp", 0)
    oldest = max(last_ts + 1 if isinstance(last_ts, int) else 0, now - int(os.environ.get("UPOW_MAX_ROLLOVERS", "3600")))
    done = _searched.get(key)
    if done is not None and done[1] <= now <= done[0]:
        # this second was searched already: nothing new until the wall clock moves (the reference would hash
        # the same 2^32 candidates again)
        raise OverflowError("int too big to convert")
    if done is not None and done[0] < now:
        oldest = max(oldest, done[0] + 1)
    max_rollovers = max(now - oldest, 0)

    ctx = get_context()
    ctx.set_template(0, prefix, rule.prev_hash, rule.n_int, rule.charset_count)
    with _ProgressLines(ctx, start_nonce, started):
        rc, res = ctx.search_min(0, start_nonce, nonce_increment, time_limit_s=STALE_AFTER_S,
                                 max_rollovers=max_rollovers, rollover_backwards=True)
    elapsed = max(time.time() - started, 1e-9)
    print(f"Worker {start_nonce + 1}: " + str(int(res.hashes_done / elapsed / 1000)) + "k hash/s")
    if rc == _lib.UPOW_EXHAUSTED and elapsed <= STALE_AFTER_S:
        # every timestamp of the window is exhausted.  The reference dies here: nonce.to_bytes(4, ...) at 2**32
        # (miner.py:87)
        lo_new = now - max_rollovers
        if done is not None and done[0] + 1 >= lo_new:
            lo_new = min(lo_new, done[1])      # joins the stretch searched earlier
        _searched[key] = (now, lo_new)
        if len(_searched) > 64:
            _searched.pop(next(iter(_searched)))
        raise OverflowError("int too big to convert")
    if rc != _lib.UPOW_OK:
        return None  # stale or cancelled: nothing is submitted (miner.py:96-98)

    candidate = with_timestamp(prefix, res.timestamp) + res.nonce.to_bytes(4, ENDIAN)
    _submit(requests, candidate.hex(), tx_hashes, last_block["id"] + 1)
    exit()


class _ProgressLines:
    """While a search runs: the reference's "Worker N: <rate>k hash/s" line (miner.py:89-95) every few seconds.
    (The reference prints it every 5e6 hashes, which is every ~5 s on a CPU core and 3800 times a second here.)"""

    def __init__(self, ctx, start_nonce: int, started: float):
        import threading
        self.ctx, self.start_nonce, self.started = ctx, start_nonce, started
        self._stop = threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.wait(PROGRESS_PERIOD_S):
            elapsed = max(time.time() - self.started, 1e-9)
            hashes = self.ctx.progress() if hasattr(self.ctx, "progress") else 0
            print(f"Worker {self.start_nonce + 1}: " + str(int(hashes / elapsed / 1000)) + "k hash/s")

    def __enter__(self):
        self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._t.join(timeout=1)


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
    last_fetch = 0.0
    while True:
        print(f"Starting {n_gpus} workers")
        info = None
        while info is None:
            wait = MINING_INFO_MIN_PERIOD_S - (time.time() - last_fetch)
            if wait > 0:
                time.sleep(wait)    # the node answers 429 above 30 requests a minute
            last_fetch = time.time()
            try:
                info = requests.get(_node_url() + "get_mining_info", timeout=5).json()["result"]
            except Exception as e:  # the reference retries forever, once a second
                print(e)
                time.sleep(1)
        if hasattr(get_context(), "clear_cancel"):
            get_context().clear_cancel()   # a cancel raised for the PREVIOUS template must not stop this one
        # optional extras beyond the reference (all off by default so the CLI behaves like miner.py):
        #   UPOW_TEMPLATES=N     race N templates of this reply (timestamps x transaction subsets), multiplier.py
        #   UPOW_STALE_POLL_S=s  abandon the search as soon as the node's tip moves (HTTP poll), stale.py
        #   UPOW_STALE_WS=url    the same through the node's /ws "block" channel, e.g. ws://localhost:3006/ws
        n_templates = int(os.environ.get("UPOW_TEMPLATES", "1"))
        poll_s = float(os.environ.get("UPOW_STALE_POLL_S", "0"))
        ws_url = os.environ.get("UPOW_STALE_WS", "")
        watcher = None
        if ws_url:
            from . import stale
            watcher = stale.WebsocketWatcher(get_context(), ws_url)
            watcher.__enter__()
        elif poll_s > 0:
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
