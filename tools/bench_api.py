#!/usr/bin/env python
"""Throughput of the drop-in PYTHON API (`QueryRouter.route_queries`) with a 1M-entry cache:
what a caller of the reference's classes gets per second once embeddings are available, and
where the time goes (host features in C++ threads, one fused device pass, Python bookkeeping)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]


class _NoEmbedder:
    """Embeddings are passed in (`q_embs=`) and the centroids are explicit, so nothing is ever encoded."""

    def encode(self, texts, **_):
        raise RuntimeError("bench_api supplies embeddings itself")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--entries", type=int, default=1_000_000)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--sections", action="store_true",
                    help="also report where a batch's time goes (wall-clock wrappers around the stages)")
    a = ap.parse_args()
    import query_router_engine as Q
    from qre_b200 import synth
    rs = np.random.Generator(np.random.PCG64(3))
    d = 384
    cents = rs.standard_normal((2, d)).astype(np.float32)
    cfg = dict(Q.PRODUCTION_CFG, cache_max_size=a.entries + a.batch * (a.rounds + 2), max_batch=a.batch,
               centroids=cents, centroid_owner=[0, 1], cache_ttl_seconds=10**6, embedder=_NoEmbedder())
    qr = Q.QueryRouter("hybrid", cfg)
    # bulk-load the device store directly (the Python insert path is per entry), then register
    # nothing on the host: these rows only serve as scan load; hits come from the batches below.
    eng = qr._engine
    n0 = a.entries
    rows = rs.standard_normal((n0, d), dtype=np.float32)
    t0 = time.perf_counter()
    eng.append_many(rows, np.full(n0, 999, np.uint64), np.zeros(n0, np.int64), np.arange(1, n0 + 1, dtype=np.uint64),
                    np.zeros(n0, np.uint8), np.ones(n0))
    load_s = time.perf_counter() - t0
    sec = {}
    if a.sections:
        import functools
        import cache as cache_mod
        from qre_b200 import _cabi

        def wrap(obj, name, label=None):
            f = getattr(obj, name)

            @functools.wraps(f)
            def g(*aa, **kk):
                t = time.perf_counter()
                try:
                    return f(*aa, **kk)
                finally:
                    sec[label or name] = sec.get(label or name, 0.0) + time.perf_counter() - t
            setattr(obj, name, g)
        wrap(cache_mod.QueryCache, "insert_batch"); wrap(cache_mod.QueryCache, "_batch_pass")
        wrap(cache_mod.QueryCache, "_apply_batch_hits"); wrap(cache_mod.QueryCache, "_flush")
        wrap(Q.QueryRouter, "_finish_batch"); wrap(_cabi, "host_features_packed")
        wrap(type(eng), "scan", "engine.scan"); wrap(type(eng), "append_many", "engine.append_many")
        wrap(type(eng), "update_many", "engine.update_many")
    words = ["alpha", "beta", "gamma", "delta", "what is", "implement", "compare", "hello", "report", "weather"]
    res = []
    for r in range(a.rounds):
        qs = [f"{words[(i + r) % len(words)]} topic {r}-{i}" for i in range(a.batch)]
        emb = rs.standard_normal((a.batch, d), dtype=np.float32)
        t0 = time.perf_counter()
        dec = qr.route_queries(qs, [None] * a.batch, ["ctx"] * a.batch, q_embs=emb)
        t1 = time.perf_counter()
        dec2 = qr.route_queries(qs, [None] * a.batch, ["ctx"] * a.batch, q_embs=emb)     # all cached now
        t2 = time.perf_counter()
        res.append(dict(miss_pass_qps=a.batch / (t1 - t0), cached_pass_qps=a.batch / (t2 - t1),
                        cached=sum(x.method.endswith("_cached") for x in dec2), routed=sum(x.method == "hybrid" for x in dec)))
    out = {"workload": f"route_queries, batch {a.batch}, {a.entries} filler rows on the device",
           "bulk_load_s": load_s, "rounds": res}
    if a.sections:
        out["ms_per_batch_by_stage_avg_of_both_passes"] = {k: round(v / (2 * a.rounds) * 1e3, 3) for k, v in sec.items()}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
