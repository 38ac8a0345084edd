"""Import the UNMODIFIED reference (`/root/reference/src`) with stubbed third-party modules.
TEST INFRASTRUCTURE ONLY — used by `oracle/gen_golden.py` and by the CPU tests that
re-validate the restatement when `/root/reference` is present (build container only; the
GPU box has no `/root/reference`, and nothing there imports this module).

Stubs (see `oracle/stubs/`): `sentence_transformers` (absent here; qre:37),
`litellm` (absent; qre:44, token_counter.py:2), `pexpect` (absent; server_manager.py:3).
"""
from __future__ import annotations

import importlib
import os
import sys
from datetime import datetime, timedelta

REFERENCE_SRC = "/root/reference/src"
STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_SRC, "cache.py"))


class FakeClock:
    """Drives `datetime.now()` inside the reference's cache.py / qre (cache.py:100,347,364,524)."""

    def __init__(self, start: datetime = datetime(2026, 1, 1, 12, 0, 0)):
        self.t = start

    def advance(self, seconds: float) -> None:
        self.t = self.t + timedelta(seconds=seconds)

    def make_datetime_class(self):
        clock = self

        class _DT(datetime):
            @classmethod
            def now(cls, tz=None):
                t = clock.t
                return cls(t.year, t.month, t.day, t.hour, t.minute, t.second, t.microsecond)

        return _DT


def load_reference(clock: FakeClock | None = None):
    """Returns (cache_module, qre_module, stub_st_module) freshly imported from the reference."""
    if not available():
        raise RuntimeError("reference not present at " + REFERENCE_SRC)
    sys.dont_write_bytecode = True
    for name in ("cache", "query_router_engine", "sentence_transformers", "litellm", "pexpect"):
        sys.modules.pop(name, None)
    saved = list(sys.path)
    sys.path.insert(0, REFERENCE_SRC)
    sys.path.insert(0, STUBS)
    try:
        st = importlib.import_module("sentence_transformers")
        assert st.__file__.startswith(STUBS)
        cache = importlib.import_module("cache")
        qre = importlib.import_module("query_router_engine")
        assert cache.__file__.startswith(REFERENCE_SRC), cache.__file__
        assert qre.__file__.startswith(REFERENCE_SRC), qre.__file__
    finally:
        sys.path[:] = saved
    if clock is not None:
        dt = clock.make_datetime_class()
        cache.datetime = dt
        qre.datetime = dt
    # leave no trace that could shadow the product's same-named modules
    for name in ("cache", "query_router_engine"):
        sys.modules.pop(name, None)
    return cache, qre, st
