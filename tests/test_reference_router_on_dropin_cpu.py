"""CPU: the reference's OWN orchestrator (`/root/reference/src/router.py`, unmodified, devices
stubbed) replayed on top of (a) the reference's cache / query_router_engine and (b) the drop-in
modules of this repo (oracle-backed engine standing in for the GPU): every result dict, token
count and chosen device must agree, for every strategy, over a growing conversation
(the loop shape of src/tests/routing_chatbot_tester.py:402-410).  Build container only."""
import os

import pytest

from oracle import ref_harness
from tests.fake_engine import OracleEngine

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _replay(dropin, strategy):
    clock = ref_harness.FakeClock()
    router_mod, cache_mod, qre_mod, _ = ref_harness.load_reference_router(
        os.path.join(ROOT, "distributed-llm_b200") if dropin else None, clock)
    if dropin:
        cache_mod.Engine = OracleEngine
        qre_mod.Engine = OracleEngine
    import sys
    sys.path.insert(0, os.path.join(ref_harness.REFERENCE_SRC, "tests"))
    try:
        import query_sets
    finally:
        sys.path.pop(0)
        sys.modules.pop("query_sets", None)
    cwd = os.getcwd()
    os.chdir("/root/reference")               # semantic_label_path is relative (qre:708)
    try:
        r = router_mod.Router(strategy=strategy)       # PRODUCTION_CFG: routing cache + response cache on
        r.nano.process = lambda hist: {"response": f"nano answer to turn {len(hist)}"}
        r.orin.process = lambda hist: {"message": {"content": f"orin answer to turn {len(hist)}"}}
        out = []
        for rnd in range(2):
            for set_name in ("general_knowledge", "personal_health"):
                history = []
                for item in query_sets.query_sets[set_name]:
                    clock.advance(2.0)
                    history.append({"role": "user", "content": item["query"]})
                    res, tokens, device = r.route_query(history)
                    res = {k: v for k, v in res.items() if k != "routing_overhead_ms"}
                    out.append((res, tokens, device))
                    history.append({"role": "assistant", "content": res["response"]})
        stats = {k: v for k, v in r.query_router.get_cache_stats().items() if k != "top_queries"}
    finally:
        os.chdir(cwd)
    return out, stats


@pytest.mark.parametrize("strategy", ["hybrid", "semantic", "token", "heuristic", "perf"])
def test_reference_router_unchanged_on_dropin_modules(strategy, capsys):
    ref_out, ref_stats = _replay(False, strategy)
    our_out, our_stats = _replay(True, strategy)
    capsys.readouterr()                                   # the reference prints its decisions
    assert len(ref_out) == len(our_out) == 44
    for i, (a, b) in enumerate(zip(ref_out, our_out)):
        assert a == b, (i, a, b)
    assert ref_stats == our_stats
    methods = {a[0]["routing_method"] for a in ref_out}
    assert "response_cache" in methods and len(methods) >= 2
