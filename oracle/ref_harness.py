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

    def make_ticking_datetime_class(self, step_us: int = 1250):
        """For the orchestrator (router.py:159,170 time a device call with two `datetime.now()`
        reads and feed the difference to the perf strategy): every read advances by a fixed step,
        so a device call always 'takes' step_us and two replays see identical latencies."""
        clock = self
        ticks = [0]

        class _TickDT(datetime):
            @classmethod
            def now(cls, tz=None):
                ticks[0] += 1
                t = clock.t + timedelta(microseconds=step_us * ticks[0])
                return cls(t.year, t.month, t.day, t.hour, t.minute, t.second, t.microsecond)

        return _TickDT


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


def load_reference_router(dropin_dir: str | None, clock: FakeClock | None = None):
    """Import the reference's ORCHESTRATOR (`src/router.py`, unmodified) together with its
    device clients, either on top of the reference's own `cache` / `query_router_engine`
    (dropin_dir=None) or on top of the drop-in modules found in `dropin_dir` — the claim of
    INTEGRATION.md §1.  Returns (router_module, cache_module, qre_module, stub_st_module)."""
    if not available():
        raise RuntimeError("reference not present at " + REFERENCE_SRC)
    sys.dont_write_bytecode = True
    names = ("router", "cache", "query_router_engine", "token_counter", "models", "models.nano", "models.orin",
             "models.server_manager", "sentence_transformers", "litellm", "pexpect")
    for name in names:
        sys.modules.pop(name, None)
    saved = list(sys.path)
    sys.path.insert(0, REFERENCE_SRC)
    if dropin_dir is not None:
        sys.path.insert(0, dropin_dir)          # shadows cache.py / query_router_engine.py only
    sys.path.insert(0, STUBS)
    try:
        st = importlib.import_module("sentence_transformers")
        router = importlib.import_module("router")
        cache = sys.modules["cache"]
        qre = sys.modules["query_router_engine"]
        assert router.__file__.startswith(REFERENCE_SRC), router.__file__
        want = dropin_dir if dropin_dir is not None else REFERENCE_SRC
        assert cache.__file__.startswith(want) and qre.__file__.startswith(want), (cache.__file__, qre.__file__)
    finally:
        sys.path[:] = saved
    if clock is not None:
        dt = clock.make_datetime_class()
        cache.datetime = dt
        qre.datetime = dt
        router.datetime = clock.make_ticking_datetime_class()   # device latency = a fixed tick, not wall clock
    for name in names:
        sys.modules.pop(name, None)
    return router, cache, qre, st
