"""CPU: `router_batch.route_queries` over the reference's unmodified orchestrator equals a loop of
`Router.route_query` (oracle-backed engine standing in for the GPU).  Build container only."""
import os
import sys

import pytest

from oracle import ref_harness
from tests.fake_engine import OracleEngine

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "distributed-llm_b200")


def _router(strategy):
    clock = ref_harness.FakeClock()
    router_mod, cache_mod, qre_mod, _ = ref_harness.load_reference_router(PKG, clock)
    cache_mod.Engine = OracleEngine
    qre_mod.Engine = OracleEngine
    cwd = os.getcwd()
    os.chdir("/root/reference")
    try:
        r = router_mod.Router(strategy=strategy)
    finally:
        os.chdir(cwd)
    r.nano.process = lambda hist: {"response": f"nano answer to turn {len(hist)}"}
    r.orin.process = lambda hist: {"message": {"content": f"orin answer to turn {len(hist)}"}}
    return r, clock


def _histories():
    sys.path.insert(0, os.path.join(ref_harness.REFERENCE_SRC, "tests"))
    try:
        import query_sets
    finally:
        sys.path.pop(0)
        sys.modules.pop("query_sets", None)
    hs = []
    for name in ("general_knowledge", "personal_health"):
        for item in query_sets.query_sets[name]:
            hs.append([{"role": "user", "content": "earlier turn about " + name},
                       {"role": "assistant", "content": "noted"},
                       {"role": "user", "content": item["query"]}])
    return hs


def _strip(res):
    d, tokens, device = res
    return {k: v for k, v in d.items() if k != "routing_overhead_ms"}, tokens, device


@pytest.mark.parametrize("strategy", ["hybrid", "semantic"])
def test_batch_equals_loop_of_route_query(strategy, capsys):
    import importlib
    sys.path.insert(0, PKG)
    try:
        rb = importlib.import_module("router_batch")
    finally:
        sys.path.pop(0)
    hs = _histories()
    a, _ = _router(strategy)
    b, _ = _router(strategy)
    for rnd in range(2):                      # second round: routing-cache and response-cache hits
        loop = [_strip(a.route_query(h)) for h in hs]
        batch = [_strip(x) for x in rb.route_queries(b, hs)]
        capsys.readouterr()
        assert len(loop) == len(batch) == len(hs)
        for i, (x, y) in enumerate(zip(loop, batch)):
            assert x == y, (rnd, i, x, y)
    sa = {k: v for k, v in a.query_router.get_cache_stats().items() if k != "top_queries"}
    sb = {k: v for k, v in b.query_router.get_cache_stats().items() if k != "top_queries"}
    assert sa == sb
    assert b.query_router is not None and type(b.query_router).__name__ == "QueryRouter"
