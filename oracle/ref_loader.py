"""Run the reference's own miner / verifier verbatim -- TEST INFRASTRUCTURE ONLY.

Works only where /root/reference exists (the build container); nothing that
runs on the GPU box imports this.  It is used by ``oracle/gen_golden.py`` to
produce the committed fixtures under ``tests/golden/`` and by
``tests/test_oracle.py`` (skipped when the reference tree is absent) to check
the restated oracle against the real code.

``miner.py`` imports ``upow.constants`` -> ``fastecdsa`` and ``upow.helpers``
-> ``base58`` / ``fastecdsa`` / ``icecream`` (upow/constants.py:1,
upow/helpers.py:11-14); none is installed here and none is used by the nonce
loop, so empty stub modules are registered before the import.  ``upow.manager``
additionally pulls ``asyncpg`` / ``pickledb`` (upow/database.py:10-12,22) and
creates ``./logs`` in the CWD (upow/my_logger.py:33-34), so it is imported from
a temp dir.
"""
from __future__ import annotations

import asyncio
import contextlib
import importlib
import io
import os
import sys
import tempfile
import types
from decimal import Decimal

REFERENCE_ROOT = os.environ.get("UPOW_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "miner.py"))


def _install_stubs():
    from . import upow_oracle as _o

    class _StubModule(types.ModuleType):
        def __getattr__(self, name):  # any other name the reference imports: an inert class
            if name.startswith("__"):
                raise AttributeError(name)
            return type(name, (Exception,), {})

    def mod(name, **attrs):
        m = sys.modules.get(name)
        if m is None:
            m = _StubModule(name)
            sys.modules[name] = m
        for k, v in attrs.items():
            setattr(m, k, v)
        return m

    class _Dummy:
        def __init__(self, *a, **k):
            pass

        def __call__(self, *a, **k):
            return None

        def __getattr__(self, name):
            return _Dummy()

    ic = _Dummy()
    ic.configureOutput = lambda **k: None
    mod("fastecdsa")
    mod("fastecdsa.curve", P256=_Dummy())
    mod("fastecdsa.point", Point=_Dummy)
    mod("fastecdsa.util", mod_sqrt=lambda *a, **k: None)
    mod("fastecdsa.keys", gen_keypair=_Dummy(), get_public_key=_Dummy())
    mod("fastecdsa.ecdsa", sign=_Dummy(), verify=_Dummy())
    sys.modules["fastecdsa"].curve = sys.modules["fastecdsa.curve"]
    sys.modules["fastecdsa"].keys = sys.modules["fastecdsa.keys"]
    sys.modules["fastecdsa"].ecdsa = sys.modules["fastecdsa.ecdsa"]
    mod("icecream", ic=ic)
    mod("base58", b58decode=_o.b58decode, b58encode=lambda b: _o.b58encode(b).encode())
    mod("asyncpg", Connection=_Dummy, Pool=_Dummy, Record=dict, create_pool=_Dummy(),
        UniqueViolationError=Exception)
    mod("pickledb", load=lambda *a, **k: _Dummy())


_miner = None
_manager = None


def load_miner(wallet: str):
    """Import /root/reference/miner.py (once) and point sys.argv[1] at ``wallet``."""
    global _miner
    if not reference_available():
        raise RuntimeError("reference tree not present at " + REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    saved = sys.argv
    sys.argv = ["miner.py", wallet, "1"]
    try:
        if _miner is None:
            _miner = importlib.import_module("miner")
    finally:
        sys.argv = saved
    return _miner


def run_mine_block(wallet: str, mining_params: dict, fixed_timestamp: int,
                   start_nonce: int = 0, nonce_increment: int = 1) -> dict:
    """``miner.mine_block(start, inc, params)`` verbatim; returns what it POSTs.

    requests.post and timestamp() are replaced; everything else is the reference's code.
    """
    miner = load_miner(wallet)
    captured = {}

    class _Resp:
        def json(self):
            return {"ok": True}

    def fake_post(url, json=None, timeout=None):
        captured.update(json)
        captured["url"] = url
        return _Resp()

    saved_argv, saved_post, saved_ts = sys.argv, miner.requests.post, miner.timestamp
    sys.argv = ["miner.py", wallet, "1"]
    miner.requests.post = fake_post
    miner.timestamp = lambda: fixed_timestamp
    params = {k: (dict(v) if isinstance(v, dict) else v) for k, v in mining_params.items()}
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            try:
                miner.mine_block(start_nonce, nonce_increment, params)
            except SystemExit:
                pass
    finally:
        sys.argv, miner.requests.post, miner.timestamp = saved_argv, saved_post, saved_ts
    return captured


def load_manager():
    global _manager
    if _manager is None:
        _install_stubs()
        if REFERENCE_ROOT not in sys.path:
            sys.path.insert(0, REFERENCE_ROOT)
        cwd = os.getcwd()
        tmp = tempfile.mkdtemp(prefix="upow_ref_")
        os.chdir(tmp)
        saved = sys.argv
        sys.argv = ["x", "--nologs"]
        try:
            _manager = importlib.import_module("upow.manager")
        finally:
            sys.argv = saved
            os.chdir(cwd)
    return _manager


def check_block_is_valid(block_content_hex: str, difficulty, last_block: dict) -> bool:
    """upow/manager.py:130-151 run verbatim."""
    mgr = load_manager()
    return bool(asyncio.run(mgr.check_block_is_valid(block_content_hex,
                                                     (Decimal(str(difficulty)), dict(last_block)))))
