"""Stale-work abort.

The reference notices a new chain tip only when its 100 s round timer fires (miner.py:149-156), so up to
100 s of work goes into a template nobody can accept any more.  Here a watcher thread asks the node for the
current tip -- by polling GET get_mining_info (rate limited to 30/minute, upow/node/main.py:675-676, hence the
default 2.5 s period) or through the node's /ws "block" channel (subscribe_block -> "new_block" messages,
websocket/socket_handlers.py:93-117, websocket/socket_utils.py:17-43) when aiohttp is importable -- and calls
upow_cancel as soon as the block being built on is no longer the last one.  upow_search_min then returns
UPOW_CANCELLED and the miner fetches a fresh template.  The cancel flag is sticky (include/upow_b200.h): a watcher
that fires BEFORE the search has started still stops it; miner.main clears the flag when it has a fresh template.
"""
from __future__ import annotations

import threading
import time
from typing import Callable, Optional


class StaleWatcher:
    """Cancels `ctx` when `fetch_tip()` reports a block id other than `building_on_id`."""

    def __init__(self, ctx, building_on_id: int, fetch_tip: Callable[[], Optional[int]], period_s: float = 2.5):
        self.ctx, self.building_on_id, self.fetch_tip, self.period_s = ctx, building_on_id, fetch_tip, period_s
        self.fired = False
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def _run(self):
        while not self._stop.wait(self.period_s):
            try:
                tip = self.fetch_tip()
            except Exception:
                continue            # the reference also just retries its HTTP calls (miner.py:131-138)
            if tip is not None and tip != self.building_on_id:
                self.fired = True
                self.ctx.cancel()
                return

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=self.period_s + 5)


def http_tip_fetcher(node_url: str, timeout: float = 5.0) -> Callable[[], Optional[int]]:
    """fetch_tip via GET <node>/get_mining_info -> result.last_block.id."""
    def fetch():
        import requests
        reply = requests.get(node_url.rstrip("/") + "/get_mining_info", timeout=timeout).json()
        return reply["result"]["last_block"].get("id")
    return fetch


def websocket_block_listener(ctx, ws_url: str, stop: threading.Event):
    """Blocks until a "new_block" message arrives on the node's /ws block channel, then cancels ctx.
    Needs aiohttp; run it in its own thread.  Returns True when it fired."""
    import asyncio

    import aiohttp

    async def listen():
        async with aiohttp.ClientSession() as session:
            async with session.ws_connect(ws_url, heartbeat=20) as ws:
                await ws.send_json({"type": "subscribe_block"})
                while not stop.is_set():
                    try:
                        msg = await ws.receive(timeout=1.0)
                    except asyncio.TimeoutError:
                        continue
                    if msg.type == aiohttp.WSMsgType.TEXT and '"new_block"' in msg.data:
                        ctx.cancel()
                        return True
                    if msg.type in (aiohttp.WSMsgType.CLOSED, aiohttp.WSMsgType.ERROR):
                        return False
        return False

    return asyncio.run(listen())


class WebsocketWatcher:
    """Runs websocket_block_listener in a thread for the lifetime of one search (context manager)."""

    def __init__(self, ctx, ws_url: str):
        self.ctx, self.ws_url = ctx, ws_url
        self.fired = False
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def _run(self):
        try:
            self.fired = bool(websocket_block_listener(self.ctx, self.ws_url, self._stop))
        except Exception:
            self.fired = False       # no websocket: the search simply runs to its own time limit

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=5)
