"""CPU: randomized operation sequences run side by side on the LIVE reference QueryCache
(/root/reference, build container only) and on the drop-in QueryCache with the oracle-backed
engine: every observable (lookup result, order, hit counts, history lengths, stats) must agree
after every step.  Exercises the host bookkeeping far beyond the scripted golden sequence:
capacity eviction with and without stale entries, TTL expiry, refresh-in-place, exact-hash
hits, invalidate by context / pattern, clear, save/load."""
import numpy as np
import pytest

from oracle import ref_harness
from tests import replay_common as R
from tests.fake_engine import OracleEngine

pytestmark = pytest.mark.skipif(not ref_harness.available(), reason="reference tree not mounted")


def _snap(qc):
    s = qc.stats()
    return ([e.query for e in qc._store.values()], [e.hit_count for e in qc._store.values()],
            [len(e.routing_history) for e in qc._store.values()],
            [e.device_used for e in qc._store.values()],
            {k: s[k] for k in ("size", "valid", "stale", "hits", "attempts", "evictions", "hybrid_fallbacks")},
            s["top_queries"])


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_random_ops_match_live_reference(seed, tmp_path):
    clock = ref_harness.FakeClock()
    ref_cache, _, _ = ref_harness.load_reference(clock)
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(seed))
    d = 64
    base = rs.standard_normal((24, d)).astype(np.float32)
    kw = dict(max_size=int(rs.integers(4, 12)), ttl_seconds=int(rs.integers(20, 80)),
              similarity_threshold=float(rs.choice([0.85, 0.6])), prediction_confidence_threshold=0.7)
    a, b = ref_cache.QueryCache(**kw), cache_mod.QueryCache(**kw)
    ctxs = ["A", "B", "C"]
    n_hits = n_miss = 0

    def vec(i, noise):
        return (base[i] * np.float32(rs.uniform(0.5, 2.0)) + noise * rs.standard_normal(d).astype(np.float32))

    for step in range(400):
        op = rs.choice(["insert", "batch", "lookup", "lookup", "lookup", "advance", "exact", "noemb",
                        "invalidate", "saveload", "clear", "warm"],
                       p=[.2, .13, .2, .15, .08, .1, .05, .03, .02, .015, .005, .02])
        if op == "insert":
            i = int(rs.integers(0, 24))
            args = (f"Query {i}", str(rs.choice(ctxs)), str(rs.choice(["nano", "orin"])))
            kwargs = dict(confidence=float(rs.uniform(0.1, 1.0)), method="hybrid", q_emb=vec(i, 0.05))
            a.insert(*args, **kwargs); b.insert(*args, **kwargs)
        elif op == "batch":
            # the additive batch form on the drop-in against the reference's loop of insert: repeats inside the
            # batch, refreshes of live keys and evictions at capacity included
            n = int(rs.integers(1, 7))
            ids = [int(rs.integers(0, 24)) for _ in range(n)]
            qs, cks = [f"Query {i}" for i in ids], [str(rs.choice(ctxs)) for _ in ids]
            devs = [str(rs.choice(["nano", "orin"])) for _ in ids]
            confs = [float(rs.uniform(0.1, 1.0)) for _ in ids]
            embs = np.stack([vec(i, 0.05) for i in ids])
            for j in range(n):
                a.insert(qs[j], cks[j], devs[j], confidence=confs[j], method="hybrid", q_emb=embs[j])
            b.insert_batch(qs, cks, devs, confs, "hybrid", embs)
        elif op == "warm":
            ids = [int(rs.integers(0, 24)) for _ in range(int(rs.integers(1, 5)))]
            table = {f"Warm {i}": vec(i, 0.05) for i in ids}

            class Emb:
                def encode(self, texts):
                    return np.stack([table[t] for t in texts])
            pairs = [(f"Warm {i}", str(rs.choice(ctxs)), str(rs.choice(["nano", "orin"]))) for i in ids]
            a.warm_up(pairs, Emb()); b.warm_up(pairs, Emb())
        elif op == "noemb":
            args = (f"Bare {int(rs.integers(0, 5))}", str(rs.choice(ctxs)), "orin")
            a.insert(*args); b.insert(*args)
        elif op in ("lookup", "exact"):
            i = int(rs.integers(0, 24))
            q = f"query {i} " if op == "exact" else f"probe {step}"
            ctx = str(rs.choice(ctxs))
            v = vec(i, float(rs.choice([0.02, 0.05, 0.3, 1.0])))
            if rs.random() < 0.05:
                v = None
            ra, rb = a.lookup(q, ctx, v), b.lookup(q, ctx, v)
            assert (ra is None) == (rb is None), (step, op, q)
            n_hits += ra is not None
            n_miss += ra is None
            if ra is not None:
                assert (ra.entry.query, ra.predicted_device, ra.predicted_confidence, ra.use_hybrid_fallback) == \
                       (rb.entry.query, rb.predicted_device, rb.predicted_confidence, rb.use_hybrid_fallback), step
        elif op == "advance":
            clock.advance(float(rs.choice([0.5, 5.0, 30.0, 90.0])))
        elif op == "invalidate":
            if rs.random() < 0.5:
                k = dict(context_key=str(rs.choice(ctxs)))
            else:
                k = dict(query_pattern=rf"query {int(rs.integers(0, 3))}\d?$")
            assert a.invalidate(**k) == b.invalidate(**k)
        elif op == "saveload":
            pa, pb = str(tmp_path / f"a{step}.json"), str(tmp_path / f"b{step}.json")
            a.save(pa); b.save(pb)
            a2, b2 = ref_cache.QueryCache(**kw), cache_mod.QueryCache(**kw)
            # each implementation loads the OTHER one's file: the schema is interchangeable
            assert a2.load(pb) == b2.load(pa)
            assert _snap(a2)[:4] == _snap(b2)[:4]
        elif op == "clear":
            a.clear(); b.clear()
        sa, sb = _snap(a), _snap(b)
        assert sa == sb, (step, op, sa, sb)
    assert n_hits > 0 and n_miss > 5 and n_hits + n_miss > 100


@pytest.mark.parametrize("seed,strategy", [(0, "hybrid"), (1, "hybrid"), (2, "semantic"), (3, "token"),
                                           (4, "heuristic"), (5, "perf")])
def test_random_route_query_matches_live_reference(seed, strategy):
    """QueryRouter.route_query driven with random queries / contexts / clock on both
    implementations: same device, method, cache_hit flag, confidence and reasoning text."""
    import zlib
    clock = ref_harness.FakeClock()
    _, ref_qre, ref_st = ref_harness.load_reference(clock)
    _, qre_mod, st = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(100 + seed))
    d = 384
    u = rs.standard_normal((2, d))
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    cents = np.stack([u[0] + 0.3 * u[1], u[1] + 0.3 * u[0]]).astype(np.float32)
    topics = ["hello there", "what is the capital of France", "implement a red black tree in python",
              "compare TCP and UDP? which is faster? why?", "tell me a joke", "summarize everything so far",
              "x == y != z;", "lorem ipsum dolor sit amet consectetur " * 4, "explain " + "very " * 60 + "long",
              "2+2?", "how often should I exercise"]
    table = {}
    pool = []
    for t in topics:
        base = rs.uniform(-0.3, 1.0) * u[0] + rs.uniform(-0.3, 1.0) * u[1] + 0.7 * rs.standard_normal(d) / np.sqrt(d)
        for v in range(3):
            text = t if v == 0 else f"{t} (variant {v})"
            table[text] = (base + (0.0 if v == 0 else rs.choice([0.03, 0.4])) * rs.standard_normal(d) / np.sqrt(d)
                           ).astype(np.float32)
            pool.append(text)
    for mod in (ref_st, st):
        mod.set_table(table, dim=d)
    cfg = dict(ref_qre.PRODUCTION_CFG)
    cfg["cache_max_size"] = 12
    cfg["cache_ttl_seconds"] = 60
    ours_cfg = dict(cfg)
    ours_cfg["centroids"], ours_cfg["centroid_owner"] = cents, [0, 1]
    a = ref_qre.QueryRouter(strategy=strategy, config=cfg)
    b = qre_mod.QueryRouter(strategy=strategy, config=ours_cfg)
    # give the reference the same centroids (its own come from the label file through the stub)
    sem_a = a.router if strategy == "semantic" else (a.router.routers["semantic"] if strategy == "hybrid" else None)
    if sem_a is not None:
        sem_a.nano_center, sem_a.orin_center = cents[0], cents[1]
    ctx_pool = [None, "user: hi", "u: " + "earlier turn. " * 100, "u: " + "long history " * 300]
    kinds = set()
    for step in range(250):
        if rs.random() < 0.15:
            clock.advance(float(rs.choice([1.0, 20.0, 70.0])))
        if strategy == "perf" and rs.random() < 0.3:
            args = (str(rs.choice(["nano", "orin"])), float(rs.uniform(50, 900)), int(rs.integers(0, 200)),
                    bool(rs.random() < 0.8))
            a.update_perf(*args); b.update_perf(*args)
        q = str(rs.choice(pool))
        context = ctx_pool[int(rs.integers(0, len(ctx_pool)))]
        key = str(rs.choice(["k1", "k2"])) if rs.random() < 0.9 else None
        da = a.route_query(q, context=context, context_key=key)
        db = b.route_query(q, context=context, context_key=key)
        assert (da.device, da.method, da.cache_hit, da.reasoning) == (db.device, db.method, db.cache_hit, db.reasoning), \
            (step, q, da, db)
        assert da.confidence == db.confidence and da.complexity_score == db.complexity_score, (step, da, db)
        kinds.add((da.method, da.cache_hit))
        sa, sb = a.get_cache_stats(), b.get_cache_stats()
        for k in ("size", "hits", "attempts", "evictions", "hybrid_fallbacks"):
            assert sa[k] == sb[k], (step, k, sa[k], sb[k])
    assert any(m.endswith("_cached") for m, _ in kinds) and any(not c for _, c in kinds)
    del zlib
