"""CPU: host logic of the drop-in classes (exact-hash path, LRU/TTL bookkeeping, eviction
order, hit/override/insert control flow, reasoning strings, heuristic regexes) replayed
against the golden fixtures with the oracle-backed fake engine standing in for the GPU."""
import os

import numpy as np
import pytest

from oracle import ref_harness
from tests import replay_common as R
from tests.fake_engine import OracleEngine


def test_cache_ops_replay_matches_reference():
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    R.replay_cache_ops(cache_mod, clock)


def test_router_replay_matches_reference():
    clock = ref_harness.FakeClock()
    _, qre_mod, st = R.load_modules(clock, OracleEngine)
    qr = R.replay_router(qre_mod, st, clock, conf_atol=0.0)
    # one fused device pass per routed query: lookup and routing never cost two passes
    assert qr._engine.calls == len(__import__("json").load(open(os.path.join(R.GOLDEN, "router_replay.json")))["steps"])


def test_route_arith_through_hybrid_router():
    _, qre_mod, st = R.load_modules(None, OracleEngine)
    R.check_route_arith(qre_mod, st)


def test_route_queries_batch_equals_sequential():
    clock = ref_harness.FakeClock()
    _, qre_mod, st = R.load_modules(clock, OracleEngine)
    R.check_route_queries_batch(qre_mod, st, clock)


def test_save_load_roundtrip_and_reference_schema(tmp_path):
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(1))
    qc = cache_mod.QueryCache(max_size=4, ttl_seconds=100)
    vec = rs.standard_normal((6, 384)).astype(np.float32)
    for i in range(4):
        qc.insert(f"q{i}", "c", "orin" if i % 2 else "nano", confidence=0.8, method="hybrid", q_emb=vec[i])
        clock.advance(1)
    p = tmp_path / "cache.json"
    qc.save(str(p))
    import json
    data = json.loads(p.read_text())
    assert set(data[0]) == {"query", "query_hash", "context_key", "embedding", "timestamp", "device_used",
                            "response_time", "hit_count", "routing_history"}
    qc2 = cache_mod.QueryCache(max_size=2, ttl_seconds=100)       # load does not enforce max_size
    assert qc2.load(str(p)) == 4 and qc2.stats()["size"] == 4
    r = qc2.lookup("other", "c", vec[2] * 2)
    assert r is not None and r.entry.query == "q2" and r.predicted_device == "nano"
    clock.advance(200)
    assert qc2.load(str(p)) == 0                                   # everything in the file is expired now
    assert qc2.load(str(tmp_path / "missing.json")) == 0
    (tmp_path / "bad.json").write_text("{not json")
    assert qc2.load(str(tmp_path / "bad.json")) == 0


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")
def test_heuristic_router_matches_live_reference():
    _, ref_qre, _ = ref_harness.load_reference()
    _, qre_mod, _ = R.load_modules(None, OracleEngine)
    cfg = dict(ref_qre.PRODUCTION_CFG)
    a, b = ref_qre.HeuristicRouter(cfg), qre_mod.HeuristicRouter(cfg)
    import json
    steps = json.load(open(os.path.join(R.GOLDEN, "router_replay.json")))["steps"]
    texts = sorted({s["query"] for s in steps}) + ["2+2?", "a* search", "SELECT x FROM y;", "", "hi", "x" * 900,
                                                   "why? how? when?", "def f(): import os", "lorem " * 30]
    for t in texts:
        for ctx in (None, "c" * 100, "c" * 4000):
            da, db = a.route(t, ctx), b.route(t, ctx)
            assert (da.device, da.confidence, da.method, da.reasoning, da.complexity_score) == \
                   (db.device, db.confidence, db.method, db.reasoning, db.complexity_score), t
    ta, tb = ref_qre.TokenBasedRouter(cfg), qre_mod.TokenBasedRouter(cfg)
    for t in ("x", "y" * 3999, "y" * 4000, "y" * 4004, "z" * 9000):
        assert ta.route(t).__dict__ == tb.route(t).__dict__
    pa, pb = ref_qre.PerformanceAwareRouter(cfg), qre_mod.PerformanceAwareRouter(cfg)
    assert pa.route("q").__dict__ == pb.route("q").__dict__
    for dev, lat, tok, ok in [("nano", 100, 10, True), ("orin", 90, 10, True), ("orin", 500, 0, False)]:
        pa.update(dev, lat, tok, ok); pb.update(dev, lat, tok, ok)
        assert pa.route("q").__dict__ == pb.route("q").__dict__
    assert ref_qre.PRODUCTION_CFG == qre_mod.PRODUCTION_CFG and ref_qre.BENCHMARK_CFG == qre_mod.BENCHMARK_CFG
    assert ref_qre.QueryRouter._default_config(None) == qre_mod.QueryRouter._default_config(None)


def test_config_embedder_works_without_sentence_transformers(monkeypatch):
    """config["embedder"] (e.g. qre_b200.embedder.B200SentenceEncoder) must give the hybrid strategy its
    semantic router and the cache its embedder even where the sentence-transformers package is not
    installed — the situation the on-GPU embedder exists for."""
    _, qre, _ = R.load_modules(engine_cls=OracleEngine)
    monkeypatch.setattr(qre, "SENTENCE_TRANSFORMERS_AVAILABLE", False)
    monkeypatch.delattr(qre, "SentenceTransformer", raising=False)

    class Emb:
        calls = 0

        def encode(self, texts, **_):
            Emb.calls += 1
            out = np.zeros((len(texts), 8), np.float32)
            for i, t in enumerate(texts):
                out[i, len(t) % 8] = 1.0
                out[i, 0] += 0.25
            return out

    cfg = dict(qre.PRODUCTION_CFG, embedder=Emb(), centroids=np.eye(2, 8, dtype=np.float32), centroid_owner=[0, 1])
    router = qre.QueryRouter("hybrid", cfg)
    assert router._fusable() and router.cache_embedder is cfg["embedder"]
    assert router.router.routers["semantic"].embedder is cfg["embedder"]
    a = router.route_query("hello there", context_key="k")
    b = router.route_query("hello there", context_key="k")
    assert a.method == "hybrid" and b.method == "hybrid_cached" and Emb.calls >= 2
    with pytest.raises(ImportError):                      # and without an embedder the reference's error stays
        qre.QueryRouter("semantic", dict(qre.PRODUCTION_CFG))
