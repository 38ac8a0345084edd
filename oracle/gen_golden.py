#!/usr/bin/env python
"""Generate tests/golden/* by running the UNMODIFIED reference (TEST INFRASTRUCTURE ONLY).

Run in the build container only (needs /root/reference):
    python oracle/gen_golden.py
The GPU box never runs this: it only reads the committed fixtures.

Fixtures:
  cache_scan.npz    real `QueryCache.insert/lookup` (cache.py:236-370) on seeded embeddings:
                    planted hits, near-threshold plants, exact duplicates (LRU tie-break),
                    other-context rows, expired rows, zero-norm rows/queries, histories.
  route_arith.npz   real `HybridRouter.route` (qre:382-414) and its three sub-routers on
                    queries whose embeddings are steered into every semantic branch.
  cache_ops.json    scripted op sequence on a small real `QueryCache` with a fake clock:
                    capacity eviction (stale first, then LRU), refresh-in-place, TTL expiry,
                    exact-hash path, invalidate, stats.
  router_replay.json real `QueryRouter(strategy="hybrid", PRODUCTION_CFG).route_query`
                    replay (miss -> insert -> cached hit -> context override / low confidence).
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "distributed-llm_b200"))

from oracle import ref_harness  # noqa: E402
from qre_b200 import synth  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
DEV = {"nano": 0, "orin": 1}


def us(dt) -> int:
    from datetime import datetime, timedelta
    return (dt - datetime(1970, 1, 1)) // timedelta(microseconds=1)


def ref_sim(q, e):
    """The expression at cache.py:283-291, evaluated by numpy exactly as the reference does."""
    norm_q = float(np.linalg.norm(q))
    norm_c = float(np.linalg.norm(e))
    return float(np.dot(q, e) / (norm_q * norm_c))


# ----------------------------------------------------------------------------------------
def gen_cache_scan():
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = ref_harness.load_reference(clock)
    N, M, d = 768, 224, 384
    TTL = 300
    w = synth.make_workload(N, M, d, n_centroids=2, seed=11, frac_planted=0.45, frac_near=0.12,
                            frac_dup=0.06, n_zero_q=2, n_zero_rows=3, frac_expired=0.0, n_ctx=4)
    ctx_names = {int(synth.ctx_id(i)): f"{int(synth.ctx_id(i)):016x}" for i in range(4)}
    # a few exact duplicate ROWS (same embedding, same context) so ties resolve by LRU order
    rs = np.random.Generator(np.random.PCG64(5))
    dup_src = rs.choice(N // 2, 24, replace=False)
    dup_dst = N // 2 + rs.choice(N // 2, 24, replace=False)
    w.rows[dup_dst] = w.rows[dup_src]
    w.row_ctx[dup_dst] = w.row_ctx[dup_src]
    # make some queries exact copies of duplicated rows
    for i, m in enumerate(range(0, 48, 4)):
        w.queries[m] = w.rows[dup_src[i]] * np.float32(1.5)
        w.q_ctx[m] = w.row_ctx[dup_src[i]]

    qc = cache_mod.QueryCache(max_size=N + 8, ttl_seconds=TTL, similarity_threshold=0.85,
                              prediction_confidence_threshold=0.70)
    insert_time = np.zeros(N, np.int64)
    # first third inserted at t0 (will be expired at lookup time), rest later
    for j in range(N):
        if j == N // 3:
            clock.advance(200)
        qc.insert(f"entry {j}", ctx_names[int(w.row_ctx[j])], "nano" if j % 3 else "orin",
                  confidence=0.5 + 0.5 * ((j * 37) % 100) / 100.0, method="hybrid",
                  q_emb=w.rows[j])
        insert_time[j] = us(clock.t)
    # extra routing records on some rows (refresh-in-place: cache.py:345-354) — note this
    # also refreshes their timestamp and moves them to the MRU end.
    hist_ops = []
    for t in range(160):
        j = int(rs.integers(N // 3, N))
        dev = "orin" if rs.random() < 0.5 else "nano"
        conf = float(rs.uniform(0.05, 1.0))
        clock.advance(0.25)
        qc.insert(f"entry {j}", ctx_names[int(w.row_ctx[j])], dev, confidence=conf,
                  method="hybrid", q_emb=w.rows[j])
        hist_ops.append((j, DEV[dev], conf, us(clock.t)))
    clock.advance(150)  # now t0+~390 s: first third (age ~390) expired, the rest (age ~190) valid

    exp_row = np.full(M, -1, np.int64)
    exp_sim = np.zeros(M, np.float64)
    exp_pred_dev = np.full(M, -1, np.int64)
    exp_pred_conf = np.zeros(M, np.float64)
    exp_fallback = np.zeros(M, np.uint8)
    now_us = np.zeros(M, np.int64)
    for m in range(M):
        clock.advance(0.01)
        now_us[m] = us(clock.t)
        r = qc.lookup(f"query {m}", ctx_names[int(w.q_ctx[m])], w.queries[m])
        if r is not None:
            j = int(r.entry.query.split()[1])
            exp_row[m] = j
            exp_sim[m] = ref_sim(w.queries[m], r.entry.embedding)
            exp_pred_dev[m] = DEV[r.predicted_device]
            exp_pred_conf[m] = r.predicted_confidence
            exp_fallback[m] = r.use_hybrid_fallback
    st = qc.stats()
    np.savez_compressed(
        os.path.join(GOLD, "cache_scan.npz"),
        rows=w.rows, row_ctx=w.row_ctx, insert_time_us=insert_time,
        hist_ops=np.array([(j, d_, t) for j, d_, _, t in hist_ops], np.int64),
        hist_conf=np.array([c for _, _, c, _ in hist_ops], np.float64),
        queries=w.queries, q_ctx=w.q_ctx, now_us=now_us, ttl_s=np.int64(TTL),
        threshold=np.float64(0.85), pct=np.float64(0.70),
        exp_row=exp_row, exp_sim=exp_sim, exp_pred_dev=exp_pred_dev,
        exp_pred_conf=exp_pred_conf, exp_fallback=exp_fallback,
        stats_hits=np.int64(st["hits"]), stats_attempts=np.int64(st["attempts"]),
        stats_evictions=np.int64(st["evictions"]), stats_size=np.int64(st["size"]),
        stats_fallbacks=np.int64(st["hybrid_fallbacks"]),
    )
    print("cache_scan: hits", int((exp_row >= 0).sum()), "of", M, "stats", {k: st[k] for k in
          ("size", "hits", "attempts", "evictions", "hybrid_fallbacks")})


# ----------------------------------------------------------------------------------------
def gen_route_arith():
    _, qre, st = ref_harness.load_reference()
    d, M = 384, 1200
    rs = np.random.Generator(np.random.PCG64(21))
    u_n, u_o = synth._unit(rs.standard_normal((2, d)))
    labels = json.load(open(os.path.join(ref_harness.REFERENCE_SRC, "tests", "semantic_labels.json")))
    table = {}
    for row in labels:
        base = u_n if row["label"].strip().lower() == "nano" else u_o
        table[row["text"].strip()] = (base + 0.5 * synth._unit(rs.standard_normal(d))).astype(np.float32)
    templates = [
        "rq {i}",                                                    # short everyday -> nano 0.75
        "what is rq {i}",                                            # simple pattern -> nano 0.90
        "implement rq {i}",                                          # complex pattern -> orin 0.92
        "rq {i} ? ? ?",                                              # multi-question -> orin 0.80
        "rq {i} == != ;",                                            # code markers -> orin 0.88
        "rq {i} " + "lorem ipsum dolor sit amet " * 8,               # no match -> token fallback
        "rq {i} " + "z" * 900,                                       # long query -> orin 0.80
    ]
    queries, contexts, Q = [], [], np.zeros((M, d), np.float32)
    for i in range(M):
        q = templates[i % len(templates)].format(i=i)
        queries.append(q)
        n_ctx_chars = int(np.exp(rs.uniform(0, np.log(12000)))) if rs.random() < 0.7 else 0
        contexts.append("c" * n_ctx_chars if n_ctx_chars else None)
        noise = synth._unit(rs.standard_normal(d))
        u = rs.random()
        if u < 0.2:        # irrelevant: both sims tiny
            v = noise + rs.uniform(-0.03, 0.03) * u_n + rs.uniform(-0.03, 0.03) * u_o
        elif u < 0.45:     # ambiguous: nearly equal pull
            a = rs.uniform(0.1, 0.9)
            v = noise + a * u_n + (a + rs.uniform(-0.04, 0.04)) * u_o
        else:
            v = noise + rs.uniform(-0.2, 1.2) * u_n + rs.uniform(-0.2, 1.2) * u_o
        Q[i] = (v * rs.uniform(0.3, 3.0)).astype(np.float32)
        table[q] = Q[i]
    if M > 10:
        Q[7] = 0.0          # zero-norm query: sims are 0.0 (qre:136-137)
        table[queries[7]] = Q[7]
    st.set_table(table, dim=d)
    cfg = dict(qre.PRODUCTION_CFG)
    cfg["semantic_label_path"] = os.path.join(ref_harness.REFERENCE_SRC, "tests", "semantic_labels.json")
    hy = qre.HybridRouter(cfg)
    sem = hy.routers["semantic"]
    cents = np.stack([sem.nano_center, sem.orin_center]).astype(np.float32)
    out = {k: [] for k in ("tok", "tok_dev", "tok_conf", "sem_dev", "sem_conf", "sem_method",
                           "sim_nano", "sim_orin", "heur_dev", "heur_conf", "heur_method",
                           "dev", "conf")}
    meth = {"semantic": 0, "semantic_fallback_irrelevant": 1, "semantic_fallback_ambiguous": 2}
    for i in range(M):
        q, c = queries[i], contexts[i]
        dt = hy.routers["token"].route(q, c)
        ds = sem.route(q, c)
        dh = hy.routers["heuristic"].route(q, c)
        dd = hy.route(q, c)
        out["tok"].append(int(dt.complexity_score))
        out["tok_dev"].append(DEV[dt.device]); out["tok_conf"].append(dt.confidence)
        out["sem_dev"].append(DEV[ds.device]); out["sem_conf"].append(ds.confidence)
        out["sem_method"].append(meth[ds.method])
        out["sim_nano"].append(sem._cosine_similarity(Q[i], sem.nano_center))
        out["sim_orin"].append(sem._cosine_similarity(Q[i], sem.orin_center))
        out["heur_dev"].append(DEV[dh.device]); out["heur_conf"].append(dh.confidence)
        out["heur_method"].append(0 if dh.method == "heuristic" else 1)
        out["dev"].append(DEV[dd.device]); out["conf"].append(dd.confidence)
    w = cfg["weights"]
    np.savez_compressed(
        os.path.join(GOLD, "route_arith.npz"),
        Q=Q, centroids=cents, queries=np.array(queries), ctx_len=np.array(
            [len(c) if c else -1 for c in contexts], np.int64),
        weights=np.array([w["token"], w["semantic"], w["heuristic"]], np.float64),
        token_threshold=np.int64(cfg["token_threshold"]),
        **{k: np.array(v) for k, v in out.items()},
    )
    sm = np.array(out["sem_method"])
    print("route_arith: sem methods", np.bincount(sm, minlength=3), "orin", int(np.sum(out["dev"])))


# ----------------------------------------------------------------------------------------
def gen_cache_ops():
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = ref_harness.load_reference(clock)
    d = 384
    rs = np.random.Generator(np.random.PCG64(31))
    base = synth._unit(rs.standard_normal((40, d))).astype(np.float32)
    qc = cache_mod.QueryCache(max_size=8, ttl_seconds=100, similarity_threshold=0.85,
                              prediction_confidence_threshold=0.70)
    ops, vecs = [], []

    def snap():
        s = qc.stats()
        return {"order": [e.query for e in qc._store.values()],
                "hit_counts": [e.hit_count for e in qc._store.values()],
                "hist_len": [len(e.routing_history) for e in qc._store.values()],
                "size": s["size"], "hits": s["hits"], "attempts": s["attempts"],
                "evictions": s["evictions"], "hybrid_fallbacks": s["hybrid_fallbacks"],
                "valid": s["valid"], "stale": s["stale"]}

    def emb(i, cos=None):
        if cos is None:
            v = base[i] * np.float32(1.0 + 0.1 * i)
        else:
            v = synth.plant(base[i], np.array(cos), rs).astype(np.float32)
        vecs.append(v)
        return len(vecs) - 1, v

    def do(op, **kw):
        rec = {"op": op, **{k: v for k, v in kw.items() if k != "vec"}}
        if op == "advance":
            clock.advance(kw["seconds"])
        elif op == "insert":
            qc.insert(kw["query"], kw["ctx"], kw["device"], confidence=kw["confidence"],
                      method="hybrid", q_emb=kw.get("vec"))
        elif op == "lookup":
            r = qc.lookup(kw["query"], kw["ctx"], kw.get("vec"))
            rec["result"] = None if r is None else {
                "entry": r.entry.query, "pred": r.predicted_device,
                "conf": r.predicted_confidence, "fallback": r.use_hybrid_fallback}
        elif op == "invalidate":
            rec["removed"] = qc.invalidate(context_key=kw.get("ctx"), query_pattern=kw.get("pattern"))
        elif op == "clear":
            qc.clear()
        rec["after"] = snap()
        ops.append(rec)

    # fill to capacity
    for i in range(8):
        vi, v = emb(i)
        do("insert", query=f"Item {i}", ctx="A" if i < 6 else "B", device="nano" if i % 2 else "orin",
           confidence=0.9, vec=v, vec_id=vi)
        do("advance", seconds=1)
    # semantic hit on Item 2 (cos 0.95) -> moves to MRU
    vi, v = emb(2, 0.95); do("lookup", query="near two", ctx="A", vec=v, vec_id=vi)
    # miss: right vector, wrong context
    vi, v = emb(2, 0.95); do("lookup", query="near two b", ctx="B", vec=v, vec_id=vi)
    # exact-hash hit (case/space-insensitive key, cache.py:520) without an embedding
    do("lookup", query="  ITEM 3 ", ctx="A")
    # capacity eviction with nothing stale -> LRU front (Item 0) goes
    vi, v = emb(8); do("insert", query="Item 8", ctx="A", device="orin", confidence=0.8, vec=v, vec_id=vi)
    # refresh in place (Item 1): history grows, moves to MRU, embedding replaced
    vi, v = emb(9); do("insert", query="item 1", ctx="A", device="orin", confidence=0.4, vec=v, vec_id=vi)
    vi, v = emb(9, 0.97); do("lookup", query="near nine", ctx="A", vec=v, vec_id=vi)
    # mixed history -> low-confidence prediction -> use_hybrid_fallback
    vi, v = emb(9); do("insert", query="item 1", ctx="A", device="nano", confidence=0.45, vec=v, vec_id=vi)
    vi, v = emb(9, 0.99); do("lookup", query="near nine again", ctx="A", vec=v, vec_id=vi)
    # below threshold -> miss
    vi, v = emb(4, 0.80); do("lookup", query="far four", ctx="A", vec=v, vec_id=vi)
    # two identical embeddings in one context: earlier-in-LRU wins, then order flips
    vi, v = emb(10); do("insert", query="Twin a", ctx="B", device="nano", confidence=0.9, vec=v, vec_id=vi)
    do("insert", query="Twin b", ctx="B", device="orin", confidence=0.9, vec=v, vec_id=vi)
    do("lookup", query="twin probe 1", ctx="B", vec=v, vec_id=vi)
    do("lookup", query="twin probe 2", ctx="B", vec=v, vec_id=vi)
    # age everything past TTL except what we refresh now
    do("advance", seconds=95)
    vi, v = emb(11); do("insert", query="Fresh 11", ctx="A", device="nano", confidence=0.9, vec=v, vec_id=vi)
    do("advance", seconds=10)
    # insert at capacity with stale entries present -> first stale in LRU order is evicted
    vi, v = emb(12); do("insert", query="Fresh 12", ctx="A", device="nano", confidence=0.9, vec=v, vec_id=vi)
    # lookup triggers _evict_expired (cache.py:250): all stale entries go, evictions counted
    vi, v = emb(11, 0.9); do("lookup", query="near eleven", ctx="A", vec=v, vec_id=vi)
    # expired exact match is deleted and falls through to semantic (cache.py:264-265)
    do("insert", query="Old exact", ctx="A", device="nano", confidence=0.9, vec=vecs[0], vec_id=0)
    do("advance", seconds=101)
    do("insert", query="Fresh 13", ctx="A", device="orin", confidence=0.9,
       vec=vecs[0] * np.float32(2.0), vec_id=0)
    do("lookup", query="old exact", ctx="A", vec=vecs[0], vec_id=0)
    # zero-norm query -> None
    vecs.append(np.zeros(d, np.float32)); do("lookup", query="zero", ctx="A", vec=vecs[-1], vec_id=len(vecs) - 1)
    # invalidate by pattern, by context, then clear
    do("invalidate", pattern="fresh 1[23]")
    do("invalidate", ctx="A")
    do("clear")
    np.save(os.path.join(GOLD, "cache_ops_vecs.npy"), np.stack(vecs))
    json.dump(ops, open(os.path.join(GOLD, "cache_ops.json"), "w"), indent=1)
    print("cache_ops:", len(ops), "ops,", len(vecs), "vectors")


# ----------------------------------------------------------------------------------------
def gen_router_replay():
    clock = ref_harness.FakeClock()
    _, qre, st = ref_harness.load_reference(clock)
    sys.path.insert(0, os.path.join(ref_harness.REFERENCE_SRC, "tests"))
    import query_sets as qs_mod
    sys.path.pop(0)
    sys.modules.pop("query_sets", None)
    d = 384
    rs = np.random.Generator(np.random.PCG64(41))
    u_n, u_o = synth._unit(rs.standard_normal((2, d)))
    labels = json.load(open(os.path.join(ref_harness.REFERENCE_SRC, "tests", "semantic_labels.json")))
    table = {}
    for row in labels:
        base = u_n if row["label"].strip().lower() == "nano" else u_o
        table[row["text"].strip()] = (base + 0.6 * synth._unit(rs.standard_normal(d))).astype(np.float32)
    queries = []
    for name, items in qs_mod.query_sets.items():
        for it in items:
            queries.append((name, it["query"], it["expected_device"]))
    for _, q, exp in queries:
        if q not in table:
            base = u_n if exp == "nano" else u_o
            table[q] = (rs.uniform(0.2, 1.0) * base + 0.8 * synth._unit(rs.standard_normal(d))).astype(np.float32)
    # paraphrases: close to an earlier query (semantic cache hit, cos ~0.93)
    para = {}
    for _, q, _ in queries[::3]:
        p = "Could you please answer this: " + q
        table[p] = synth.plant(table[q], np.array(0.93), rs).astype(np.float32)
        para[p] = q
    st.set_table(table, dim=d)
    cfg = dict(qre.PRODUCTION_CFG)
    cfg["semantic_label_path"] = os.path.join(ref_harness.REFERENCE_SRC, "tests", "semantic_labels.json")
    qr = qre.QueryRouter(strategy="hybrid", config=cfg)
    steps = []

    def call(q, context, ctxk):
        clock.advance(1.0)
        dd = qr.route_query(q, context=context, context_key=ctxk)
        steps.append({"query": q, "context_len": len(context) if context else -1, "ctx": ctxk,
                      "device": dd.device, "confidence": dd.confidence, "method": dd.method,
                      "cache_hit": dd.cache_hit, "reasoning": dd.reasoning})

    short_ctx = "user: hi\nassistant: hello"
    long_ctx = "user: " + "tell me more about that topic please. " * 100     # 3806 chars >= 3200
    # pass 1: all misses -> inserted; pass 2: cached; pass 3: paraphrases (semantic hits);
    # pass 4: same queries under a heavy context (override of cached nano, qre:584-587);
    # pass 5: other context key (miss again).
    for _, q, _ in queries:
        call(q, short_ctx, "ctx0")
    for _, q, _ in queries:
        call(q, short_ctx, "ctx0")
    for p in para:
        call(p, short_ctx, "ctx0")
    for _, q, _ in queries[:12]:
        call(q, long_ctx, "ctx0")
    for _, q, _ in queries[:12]:
        call(q, long_ctx, "ctx0")
    for _, q, _ in queries[:6]:
        call(q, None, None)
    allq = sorted({s_["query"] for s_ in steps})
    sem = qr.router.routers["semantic"]
    np.savez_compressed(os.path.join(GOLD, "router_replay_table.npz"),
                        texts=np.array(allq), vecs=np.stack([table[t] for t in allq]),
                        centroids=np.stack([sem.nano_center, sem.orin_center]).astype(np.float32))
    json.dump({"steps": steps, "stats": {k: v for k, v in qr.get_cache_stats().items()
                                         if k != "top_queries"},
               "short_ctx": short_ctx, "long_ctx": long_ctx},
              open(os.path.join(GOLD, "router_replay.json"), "w"), indent=1)
    from collections import Counter
    print("router_replay:", len(steps), "steps", Counter(s["method"] for s in steps))




# ----------------------------------------------------------------------------------------
def config1_inputs():
    """Seeded inputs of BASELINE.json configs[0] (1k queries x 10k cache entries x 384-d fp32, hybrid
    strategy + semantic cache).  Shared by the generator and the GPU test: only the reference's
    OUTPUTS are committed, the inputs are regenerated from the seed (and checksummed)."""
    w = synth.make_workload(10_000, 1_000, 384, n_centroids=2, seed=101, frac_planted=0.45, frac_near=0.03,
                            frac_dup=0.01, n_zero_q=2, n_zero_rows=2, frac_expired=0.0, n_ctx=6)
    rs = np.random.Generator(np.random.PCG64(102))
    templates = ["note {i}", "what is item {i}", "implement step {i}", "item {i}? really? sure?",
                 "item {i} " + "lorem ipsum dolor sit amet " * 8]
    q_text = [templates[int(rs.integers(0, len(templates)))].format(i=m) for m in range(1000)]
    ctx_len = np.where(rs.random(1000) < 0.25, rs.integers(3200, 9000, 1000),
                       np.where(rs.random(1000) < 0.5, rs.integers(1, 600, 1000), 0))
    row_dev = rs.integers(0, 2, 10_000)
    row_conf = rs.uniform(0.3, 1.0, 10_000)
    # a second routing record on 30 % of the rows so predictions are not all unanimous
    second = rs.random(10_000) < 0.3
    second_dev = rs.integers(0, 2, 10_000)
    second_conf = rs.uniform(0.3, 1.0, 10_000)
    checksum = float(np.float64(w.rows.astype(np.float64).sum() + w.queries.astype(np.float64).sum()))
    return dict(w=w, q_text=q_text, ctx_len=ctx_len, row_dev=row_dev, row_conf=row_conf, second=second,
                second_dev=second_dev, second_conf=second_conf, checksum=checksum)


def run_config1(qre_mod, st_mod, clock, patch_centroids):
    """Drive one implementation (reference or drop-in) through config 1; returns the decisions."""
    I = config1_inputs()
    w = I["w"]
    names = ["nano", "orin"]
    table = {f"entry {j}": w.rows[j] for j in range(10_000)}
    table.update({I["q_text"][m]: w.queries[m] for m in range(1_000)})
    st_mod.set_table(table, dim=384)
    cfg = dict(qre_mod.PRODUCTION_CFG, cache_max_size=20_000, cache_ttl_seconds=10**6)
    cfg.update(patch_centroids(cfg))
    qr = qre_mod.QueryRouter(strategy="hybrid", config=cfg)
    patch_centroids(cfg, qr)
    ctx_key = {int(synth.ctx_id(i)): f"{int(synth.ctx_id(i)):016x}" for i in range(6)}
    for j in range(10_000):
        qr._cache.insert(f"entry {j}", ctx_key[int(w.row_ctx[j])], names[int(I["row_dev"][j])],
                         confidence=float(I["row_conf"][j]), method="hybrid", q_emb=w.rows[j])
        if I["second"][j]:
            qr._cache.insert(f"entry {j}", ctx_key[int(w.row_ctx[j])], names[int(I["second_dev"][j])],
                             confidence=float(I["second_conf"][j]), method="hybrid", q_emb=w.rows[j])
    out = []
    for m in range(1_000):
        clock.advance(0.5)
        n = int(I["ctx_len"][m])
        d = qr.route_query(I["q_text"][m], context=("c" * n) if n else None, context_key=ctx_key[int(w.q_ctx[m])])
        out.append(dict(device=d.device, confidence=d.confidence, method=d.method, cache_hit=d.cache_hit))
    st = {k: v for k, v in qr.get_cache_stats().items() if k != "top_queries"}
    return out, st, I["checksum"]


def gen_config1():
    clock = ref_harness.FakeClock()
    _, qre, st = ref_harness.load_reference(clock)
    rs = np.random.Generator(np.random.PCG64(103))
    u = synth._unit(rs.standard_normal((2, 384)))
    cents = np.stack([u[0] + 0.3 * u[1], u[1] + 0.3 * u[0]]).astype(np.float32)

    def patch(cfg, qr=None):
        if qr is not None:          # the reference builds centroids from the label file: overwrite them
            sem = qr.router.routers["semantic"]
            sem.nano_center, sem.orin_center = cents[0], cents[1]
        return {"semantic_label_path": os.path.join(ref_harness.REFERENCE_SRC, "tests", "semantic_labels.json")}

    import time
    t0 = time.time()
    out, stats, checksum = run_config1(qre, st, clock, patch)
    np.save(os.path.join(GOLD, "config1_centroids.npy"), cents)
    json.dump({"decisions": out, "stats": stats, "input_checksum": checksum,
               "reference_seconds": round(time.time() - t0, 1)},
              open(os.path.join(GOLD, "config1_replay.json"), "w"))
    from collections import Counter
    print("config1:", Counter((o["method"], o["cache_hit"]) for o in out), stats, f"{time.time() - t0:.0f}s")


def gen_embedder():
    """Embedder fixtures from the THIRD-PARTY code the reference calls (transformers' BertModel and
    BertTokenizer; sentence-transformers' Pooling(mean) + Normalize restated in three lines), not from the
    restatement: seeded random weights of two shapes (`random_weights` of oracle/minilm_oracle.py regenerates
    them from the seed stored here), token ids, expected embeddings; and the ids BertTokenizer assigns to the
    texts of tests/wordpiece_cases.py."""
    import tempfile
    import torch
    from transformers import BertConfig, BertModel, BertTokenizer
    from oracle import minilm_oracle as MO
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from wordpiece_cases import TEXTS, VOCAB
    out = {}
    shapes = {"hd32": MO.EncoderConfig(vocab_size=500, hidden=128, layers=2, heads=4, intermediate=256,
                                       max_positions=64, max_seq_len=48),
              "hd64": MO.EncoderConfig(vocab_size=400, hidden=128, layers=2, heads=2, intermediate=320,
                                       max_positions=40, max_seq_len=40),
              "h384": MO.EncoderConfig(vocab_size=600, hidden=384, layers=2, heads=12, intermediate=1536,
                                       max_positions=64, max_seq_len=64)}
    for name, cfg in shapes.items():
        seed = {"hd32": 31, "hd64": 32, "h384": 33}[name]
        w = MO.random_weights(cfg, seed=seed)
        hc = BertConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden, num_hidden_layers=cfg.layers,
                        num_attention_heads=cfg.heads, intermediate_size=cfg.intermediate,
                        max_position_embeddings=cfg.max_positions, hidden_act="gelu", layer_norm_eps=cfg.ln_eps)
        model = BertModel(hc, add_pooling_layer=False).eval()
        missing, unexpected = model.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}, strict=False)
        assert not unexpected
        rs = np.random.Generator(np.random.PCG64(seed + 100))
        lens = [1, 2, 7, 15, 16, 17, 31, 32, 33, cfg.max_seq_len, 5, 24]
        ids = [rs.integers(0, cfg.vocab_size, n).astype(np.int32) for n in lens]
        L = max(lens)
        tok = torch.zeros((len(ids), L), dtype=torch.long)
        mask = torch.zeros((len(ids), L), dtype=torch.long)
        for i, s_ in enumerate(ids):
            tok[i, :len(s_)] = torch.from_numpy(s_.astype(np.int64))
            mask[i, :len(s_)] = 1
        with torch.no_grad():
            h = model(input_ids=tok, attention_mask=mask).last_hidden_state
        mf = mask.unsqueeze(-1).float()
        pooled = (h * mf).sum(1) / mf.sum(1).clamp(min=1e-9)
        out[name + "_cfg"] = np.array([cfg.vocab_size, cfg.hidden, cfg.layers, cfg.heads, cfg.intermediate,
                                       cfg.max_positions, cfg.max_seq_len, seed], np.int64)
        out[name + "_ids"] = np.concatenate(ids)
        out[name + "_lens"] = np.array(lens, np.int32)
        out[name + "_pooled"] = pooled.numpy().astype(np.float32)
        out[name + "_emb"] = torch.nn.functional.normalize(pooled, p=2, dim=1).numpy().astype(np.float32)
    d = tempfile.mkdtemp()
    vp = os.path.join(d, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    bt = BertTokenizer(vp, do_lower_case=True)
    tok_ids = {str(ml): [bt(t, truncation=True, max_length=ml)["input_ids"] for t in TEXTS] for ml in (256, 16)}
    np.savez_compressed(os.path.join(GOLD, "embedder.npz"), **out)
    json.dump({"texts": TEXTS, "ids": tok_ids, "transformers": __import__("transformers").__version__,
               "torch": torch.__version__}, open(os.path.join(GOLD, "wordpiece.json"), "w"), ensure_ascii=True)
    print("embedder:", {k: v.shape for k, v in out.items() if k.endswith("_emb")})


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    only = sys.argv[1:] 
    gens = dict(cache_scan=gen_cache_scan, route_arith=gen_route_arith, cache_ops=gen_cache_ops,
                router_replay=gen_router_replay, config1=gen_config1, embedder=gen_embedder)
    for name, fn in gens.items():
        if not only or name in only:
            fn()
