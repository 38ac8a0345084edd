"""Replays of the golden fixtures through the drop-in host classes.  Shared by the CPU suite
(oracle-backed fake engine: host logic only) and the GPU suite (real libqre_b200 engine)."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
STUBS = os.path.join(ROOT, "oracle", "stubs")


def load_modules(clock=None, engine_cls=None):
    """Import the product's cache.py / query_router_engine.py with the stub embedder on the path
    (sentence-transformers is not installed anywhere in this project), optionally with a fake
    clock and a substitute engine class."""
    if STUBS not in sys.path:
        sys.path.insert(0, STUBS)
    for name in ("cache", "query_router_engine"):
        sys.modules.pop(name, None)
    cache = importlib.import_module("cache")
    qre = importlib.import_module("query_router_engine")
    pkg = os.path.join(ROOT, "distributed-llm_b200")
    assert cache.__file__.startswith(pkg) and qre.__file__.startswith(pkg)
    st = importlib.import_module("sentence_transformers")
    if clock is not None:
        dt = clock.make_datetime_class()
        cache.datetime = dt
        qre.datetime = dt
    if engine_cls is not None:
        cache.Engine = engine_cls
        qre.Engine = engine_cls
    return cache, qre, st


def replay_cache_ops(cache_mod, clock, rel=1e-5):
    ops = json.load(open(os.path.join(GOLDEN, "cache_ops.json")))
    vecs = np.load(os.path.join(GOLDEN, "cache_ops_vecs.npy"))
    qc = cache_mod.QueryCache(max_size=8, ttl_seconds=100, similarity_threshold=0.85,
                              prediction_confidence_threshold=0.70)
    for n, op in enumerate(ops):
        kind = op["op"]
        vec = vecs[op["vec_id"]] if "vec_id" in op else None
        if kind == "advance":
            clock.advance(op["seconds"])
        elif kind == "insert":
            qc.insert(op["query"], op["ctx"], op["device"], confidence=op["confidence"], method="hybrid",
                      q_emb=vec)
        elif kind == "lookup":
            r = qc.lookup(op["query"], op["ctx"], vec)
            exp = op["result"]
            assert (r is None) == (exp is None), (n, op["query"], r, exp)
            if r is not None:
                assert r.entry.query == exp["entry"], (n, op["query"])
                assert r.predicted_device == exp["pred"]
                assert r.predicted_confidence == exp["conf"]
                assert r.use_hybrid_fallback == exp["fallback"]
        elif kind == "invalidate":
            assert qc.invalidate(context_key=op.get("ctx"), query_pattern=op.get("pattern")) == op["removed"]
        elif kind == "clear":
            qc.clear()
        s, a = qc.stats(), op["after"]
        got = {"order": [e.query for e in qc._store.values()],
               "hit_counts": [e.hit_count for e in qc._store.values()],
               "hist_len": [len(e.routing_history) for e in qc._store.values()]}
        for k in ("order", "hit_counts", "hist_len"):
            assert got[k] == a[k], (n, kind, k, got[k], a[k])
        for k in ("size", "hits", "attempts", "evictions", "hybrid_fallbacks", "valid", "stale"):
            assert s[k] == a[k], (n, kind, k, s[k], a[k])
    return qc


def replay_router(qre_mod, st_mod, clock, conf_atol=4e-6):
    g = json.load(open(os.path.join(GOLDEN, "router_replay.json")))
    t = np.load(os.path.join(GOLDEN, "router_replay_table.npz"))
    st_mod.set_table({str(k): v for k, v in zip(t["texts"], t["vecs"])}, dim=t["vecs"].shape[1])
    cfg = dict(qre_mod.PRODUCTION_CFG)
    cfg["centroids"] = t["centroids"]
    cfg["centroid_owner"] = [0, 1]
    qr = qre_mod.QueryRouter(strategy="hybrid", config=cfg)
    ctxs = {len(g["short_ctx"]): g["short_ctx"], len(g["long_ctx"]): g["long_ctx"], -1: None}
    for n, s in enumerate(g["steps"]):
        clock.advance(1.0)
        d = qr.route_query(s["query"], context=ctxs[s["context_len"]], context_key=s["ctx"])
        assert d.device == s["device"], (n, s["query"], d, s)
        assert d.method == s["method"] and d.cache_hit == s["cache_hit"], (n, d, s)
        assert abs(d.confidence - s["confidence"]) <= conf_atol + 1e-5 * abs(s["confidence"]), (n, d, s)
        # reasoning strings carry 2-3 decimals of the same numbers: identical text
        assert d.reasoning == s["reasoning"], (n, d.reasoning, s["reasoning"])
    stats = qr.get_cache_stats()
    for k, v in g["stats"].items():
        assert stats[k] == v, (k, stats[k], v)
    return qr


def check_route_arith(qre_mod, st_mod, rel=1e-5):
    """Our HybridRouter (device epilogue) against the recorded reference HybridRouter outputs."""
    g = np.load(os.path.join(GOLDEN, "route_arith.npz"))
    Q = g["Q"]
    st_mod.set_table({str(q): v for q, v in zip(g["queries"], Q)}, dim=Q.shape[1])
    cfg = dict(qre_mod.PRODUCTION_CFG)
    cfg["centroids"] = g["centroids"]
    cfg["centroid_owner"] = [0, 1]
    hy = qre_mod.HybridRouter(cfg)
    names = ["nano", "orin"]
    sem_names = ("semantic", "semantic_fallback_irrelevant", "semantic_fallback_ambiguous")
    margin = np.abs(g["sim_orin"] - g["sim_nano"])
    band = (np.abs(margin - 0.03) < 2e-6) | (np.abs(np.maximum(g["sim_orin"], g["sim_nano"]) - 0.05) < 2e-6) \
        | (margin < 2e-6)
    n_checked = 0
    for m in range(0, Q.shape[0], 3):
        if band[m]:
            continue
        q = str(g["queries"][m])
        ctx_len = int(g["ctx_len"][m])
        context = None if ctx_len < 0 else "c" * ctx_len
        dh = hy.routers["heuristic"].route(q, context)
        assert (names.index(dh.device), dh.confidence) == (int(g["heur_dev"][m]), float(g["heur_conf"][m]))
        ds = hy.routers["semantic"].route(q, context)
        assert ds.method == sem_names[int(g["sem_method"][m])], (m, ds)
        assert names.index(ds.device) == int(g["sem_dev"][m])
        assert abs(ds.confidence - float(g["sem_conf"][m])) <= 2e-6 + rel * abs(float(g["sem_conf"][m]))
        d = hy.route(q, context)
        assert names.index(d.device) == int(g["dev"][m]), (m, d)
        assert abs(d.confidence - float(g["conf"][m])) <= 4e-6 + rel * abs(float(g["conf"][m])), (m, d)
        n_checked += 1
    assert n_checked > 300


def check_route_queries_batch(qre_mod, st_mod, clock):
    """`route_queries` (one fused device pass per batch) against a twin router driven one query at
    a time, on batches without intra-batch interactions: identical decisions, methods, strings."""
    g = json.load(open(os.path.join(GOLDEN, "router_replay.json")))
    t = np.load(os.path.join(GOLDEN, "router_replay_table.npz"))
    st_mod.set_table({str(k): v for k, v in zip(t["texts"], t["vecs"])}, dim=t["vecs"].shape[1])
    cfg = dict(qre_mod.PRODUCTION_CFG)
    cfg["centroids"], cfg["centroid_owner"] = t["centroids"], [0, 1]
    a, b = qre_mod.QueryRouter("hybrid", cfg), qre_mod.QueryRouter("hybrid", cfg)
    steps = g["steps"]
    first = [s for s in steps[:32]]                       # pass 1 of the fixture: distinct queries
    qs = [s["query"] for s in first]
    ctxs = {len(g["short_ctx"]): g["short_ctx"], len(g["long_ctx"]): g["long_ctx"], -1: None}
    for rnd, context in enumerate((g["short_ctx"], g["short_ctx"], g["long_ctx"])):
        clock.advance(5.0)
        seq = [a.route_query(q, context=context, context_key="ctx0") for q in qs]
        bat = b.route_queries(qs, [context] * len(qs), ["ctx0"] * len(qs))
        for x, y in zip(seq, bat):
            assert (x.device, x.method, x.cache_hit, x.reasoning) == (y.device, y.method, y.cache_hit, y.reasoning), \
                (rnd, x, y)
            assert x.confidence == y.confidence
    sa, sb = a.get_cache_stats(), b.get_cache_stats()
    for k in ("size", "hits", "attempts", "evictions", "hybrid_fallbacks"):
        assert sa[k] == sb[k], (k, sa[k], sb[k])
    assert {d.method for d in bat} == {"hybrid", "hybrid_cached"}
