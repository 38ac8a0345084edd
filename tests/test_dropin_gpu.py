"""GPU: the drop-in classes (cache.QueryCache, query_router_engine.*) on the real libqre_b200
engine against outputs recorded from the reference's own code (tests/golden)."""
import os

import numpy as np
import pytest

from oracle import ref_harness
from tests import replay_common as R

pytestmark = pytest.mark.gpu


def _two_gpus():
    """GPU indices for a two-shard store: two different GPUs when the box has them, else two shards
    on GPU 0 (the same code path: per-shard scans, lists gathered on the first shard, one merge)."""
    import torch
    return [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]


def test_cache_ops_replay_gpu():
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock)
    qc = R.replay_cache_ops(cache_mod, clock)
    assert type(qc._engine).__module__ == "qre_b200.engine"


def test_router_replay_gpu():
    clock = ref_harness.FakeClock()
    _, qre_mod, st = R.load_modules(clock)
    qr = R.replay_router(qre_mod, st, clock)
    assert qr._engine.shard is not None and qr._engine.shard.launch_count() > 300


def test_route_arith_gpu():
    _, qre_mod, st = R.load_modules(None)
    R.check_route_arith(qre_mod, st)


def test_route_queries_batch_gpu():
    clock = ref_harness.FakeClock()
    _, qre_mod, st = R.load_modules(clock)
    R.check_route_queries_batch(qre_mod, st, clock)


def test_golden_cache_scan_through_query_cache():
    """The recorded reference run of cache_scan.npz replayed through OUR QueryCache on the GPU:
    same inserts, same refreshes, same lookups at the same (fake) times."""
    g = np.load(os.path.join(R.GOLDEN, "cache_scan.npz"))
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock)
    from datetime import datetime, timedelta
    epoch = datetime(1970, 1, 1)

    def set_time(us):
        clock.t = epoch + timedelta(microseconds=int(us))

    N, M = g["rows"].shape[0], g["queries"].shape[0]
    qc = cache_mod.QueryCache(max_size=N + 8, ttl_seconds=int(g["ttl_s"]), similarity_threshold=0.85,
                              prediction_confidence_threshold=0.70)
    names = {int(c): f"{int(c):016x}" for c in np.unique(np.concatenate([g["row_ctx"], g["q_ctx"]]))}
    for j in range(N):
        set_time(g["insert_time_us"][j])
        qc.insert(f"entry {j}", names[int(g["row_ctx"][j])], "nano" if j % 3 else "orin",
                  confidence=0.5 + 0.5 * ((j * 37) % 100) / 100.0, method="hybrid", q_emb=g["rows"][j])
    for (j, dev, t), conf in zip(g["hist_ops"], g["hist_conf"]):
        set_time(t)
        qc.insert(f"entry {j}", names[int(g["row_ctx"][j])], "orin" if dev else "nano", confidence=float(conf),
                  method="hybrid", q_emb=g["rows"][j])
    for m in range(M):
        set_time(g["now_us"][m])
        r = qc.lookup(f"query {m}", names[int(g["q_ctx"][m])], g["queries"][m])
        exp = int(g["exp_row"][m])
        assert (r is None) == (exp < 0), m
        if r is not None:
            assert r.entry.query == f"entry {exp}", (m, r.entry.query, exp)
            assert ["nano", "orin"].index(r.predicted_device) == int(g["exp_pred_dev"][m])
            assert r.predicted_confidence == float(g["exp_pred_conf"][m])
            assert r.use_hybrid_fallback == bool(g["exp_fallback"][m])
    s = qc.stats()
    assert s["hits"] == int(g["stats_hits"]) and s["evictions"] == int(g["stats_evictions"])
    assert s["size"] == int(g["stats_size"]) and s["hybrid_fallbacks"] == int(g["stats_fallbacks"])


def test_lookup_batch_equals_sequential_snapshot():
    cache_mod, _, _ = R.load_modules(None)
    from qre_b200 import synth
    w = synth.make_workload(400, 60, 384, n_centroids=2, seed=21, n_ctx=3, frac_expired=0.0, frac_dup=0.0)
    names = {int(synth.ctx_id(i)): f"ctx{i}" for i in range(3)}
    a = cache_mod.QueryCache(max_size=500, ttl_seconds=10**6)
    b = cache_mod.QueryCache(max_size=500, ttl_seconds=10**6)
    for qc in (a, b):
        for j in range(400):
            qc.insert(f"e{j}", names[int(w.row_ctx[j])], "orin" if j % 2 else "nano", q_emb=w.rows[j])
    seq = [a.lookup(f"q{m}", names[int(w.q_ctx[m])], w.queries[m]) for m in range(60)]
    bat = b.lookup_batch([f"q{m}" for m in range(60)], [names[int(c)] for c in w.q_ctx], w.queries)
    assert [None if r is None else r.entry.query for r in seq] == [None if r is None else r.entry.query for r in bat]
    assert sum(r is not None for r in seq) > 10


def test_binary_snapshot_roundtrip(tmp_path):
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock)
    from qre_b200 import synth
    w = synth.make_workload(600, 80, 384, n_centroids=2, seed=23, n_ctx=3, frac_expired=0.0, frac_dup=0.0)
    names = {int(synth.ctx_id(i)): f"ctx{i}" for i in range(3)}
    a = cache_mod.QueryCache(max_size=700, ttl_seconds=1000)
    for j in range(600):
        a.insert(f"e{j}", names[int(w.row_ctx[j])], "orin" if j % 2 else "nano", confidence=0.9, q_emb=w.rows[j])
        if j == 299:
            clock.advance(600)                      # the first 300 will be expired at load time
    a.invalidate(query_pattern=r"^e3[0-4]\d$")       # tombstoned slots travel through the snapshot
    p = str(tmp_path / "snap.json")
    a.save_snapshot(p)
    clock.advance(500)
    b = cache_mod.QueryCache(max_size=700, ttl_seconds=1000)
    n = b.load_snapshot(p)
    assert n == 250 and b.stats()["size"] == 250
    ra = [a.lookup(f"q{m}", names[int(w.q_ctx[m])], w.queries[m]) for m in range(80)]
    rb = [b.lookup(f"q{m}", names[int(w.q_ctx[m])], w.queries[m]) for m in range(80)]
    assert [None if r is None else r.entry.query for r in ra] == [None if r is None else r.entry.query for r in rb]
    assert sum(r is not None for r in rb) > 5
    e = next(iter(b._store.values()))
    assert np.array_equal(e.embedding, w.rows[int(e.query[1:])])
    b.insert("fresh", "ctx0", "nano", q_emb=w.rows[0] * 2)      # freed slots are reusable after load
    assert b.lookup("probe", "ctx0", w.rows[0]).entry.query == "fresh"


@pytest.mark.parametrize("n_shards", [1, 2])
def test_config1_sequential_replay_matches_reference_run(n_shards):
    """BASELINE.json configs[0] — the reference's own CPU-runnable case: hybrid strategy + semantic
    cache, 10,000 pre-inserted entries, 1,000 queries routed ONE AT A TIME through
    QueryRouter.route_query (so query i can hit what query i-1 inserted), 384-d fp32.  The
    decisions recorded from the reference's own code (82 s on the build host) must be reproduced
    by the drop-in classes on the GPU: device, method and cache_hit identical, confidence within
    the score tolerance."""
    import json
    import time
    from oracle import gen_golden as G
    g = json.load(open(os.path.join(R.GOLDEN, "config1_replay.json")))
    cents = np.load(os.path.join(R.GOLDEN, "config1_centroids.npy"))
    clock = ref_harness.FakeClock()
    _, qre_mod, st = R.load_modules(clock)

    def patch(cfg, qr=None):
        extra = {"centroids": cents, "centroid_owner": [0, 1]}
        if n_shards == 2:            # config["gpus"]: rows sharded over two device stores, same decisions
            extra["gpus"] = _two_gpus()
        return extra

    t0 = time.time()
    out, stats, checksum = G.run_config1(qre_mod, st, clock, patch)
    ours_s = time.time() - t0
    assert checksum == g["input_checksum"], "seeded inputs differ from the ones the fixture was recorded on"
    bad = [i for i, (a, b) in enumerate(zip(out, g["decisions"]))
           if (a["device"], a["method"], a["cache_hit"]) != (b["device"], b["method"], b["cache_hit"])]
    assert not bad, (bad[:5], [(out[i], g["decisions"][i]) for i in bad[:3]])
    for a, b in zip(out, g["decisions"]):
        assert abs(a["confidence"] - b["confidence"]) <= 4e-6 + 1e-5 * abs(b["confidence"]), (a, b)
    for k, v in g["stats"].items():
        assert stats[k] == v, (k, stats[k], v)
    print(f"config 1 ({n_shards} shard(s)): reference {g['reference_seconds']} s, drop-in on GPU {ours_s:.1f} s")


@pytest.mark.parametrize("seed,sharded", [(0, False), (1, False), (2, True), (3, True)])
def test_random_ops_device_engine_vs_oracle_engine(seed, sharded, tmp_path):
    """Differential test of the device store under realistic interleavings: the same random
    sequence of inserts / refreshes / lookups / evictions / invalidations / snapshot round trips is
    applied to two QueryCache objects, one on the real libqre_b200 engine and one on the
    oracle-backed engine; every lookup result and the store order must agree at every step."""
    from tests.fake_engine import OracleEngine
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock)
    rs = np.random.Generator(np.random.PCG64(500 + seed))
    d = 384
    base = rs.standard_normal((40, d)).astype(np.float32)
    kw = dict(max_size=int(rs.integers(8, 20)), ttl_seconds=int(rs.integers(30, 90)), similarity_threshold=0.85,
              prediction_confidence_threshold=0.7)
    gkw = dict(kw, gpus=_two_gpus()) if sharded else kw
    gpu = cache_mod.QueryCache(**gkw)
    assert type(gpu._engine).__name__ == ("ShardedEngine" if sharded else "Engine")
    cpu = cache_mod.QueryCache(**kw, engine=OracleEngine(capacity=64, dim=d))
    ctxs = ["A", "B"]
    hits = 0
    for step in range(350):
        op = rs.choice(["insert", "lookup", "batch", "advance", "invalidate", "snapshot"],
                       p=[.35, .35, .1, .12, .05, .03])
        if op == "insert":
            i = int(rs.integers(0, 40))
            v = base[i] * np.float32(rs.uniform(0.5, 2)) + np.float32(0.03) * rs.standard_normal(d).astype(np.float32)
            args = (f"Item {i}", str(rs.choice(ctxs)), str(rs.choice(["nano", "orin"])))
            k = dict(confidence=float(rs.uniform(0.2, 1)), method="hybrid", q_emb=v)
            gpu.insert(*args, **k); cpu.insert(*args, **k)
        elif op == "lookup":
            i, ctx = int(rs.integers(0, 40)), str(rs.choice(ctxs))
            live = list(cpu._store.values())
            if live and rs.random() < 0.7:              # aim at something that is in the cache
                e = live[int(rs.integers(0, len(live)))]
                i, ctx = int(e.query.split()[1]), e.context_key
            v = base[i] + np.float32(rs.choice([0.02, 0.2, 0.8])) * rs.standard_normal(d).astype(np.float32)
            a, b = gpu.lookup(f"probe {step}", ctx, v), cpu.lookup(f"probe {step}", ctx, v)
            assert (a is None) == (b is None), step
            if a is not None:
                hits += 1
                assert (a.entry.query, a.predicted_device, a.predicted_confidence) == \
                       (b.entry.query, b.predicted_device, b.predicted_confidence), step
        elif op == "batch":
            idx = rs.integers(0, 40, 6)
            V = base[idx] + np.float32(0.05) * rs.standard_normal((6, d)).astype(np.float32)
            qs, cs = [f"bq {step}-{j}" for j in range(6)], [str(rs.choice(ctxs)) for _ in range(6)]
            a, b = gpu.lookup_batch(qs, cs, V), cpu.lookup_batch(qs, cs, V)
            assert [None if x is None else x.entry.query for x in a] == [None if x is None else x.entry.query for x in b]
        elif op == "advance":
            clock.advance(float(rs.choice([1.0, 10.0, 40.0])))
        elif op == "invalidate":
            pat = rf"item {int(rs.integers(0, 4))}\d?$"
            assert gpu.invalidate(query_pattern=pat) == cpu.invalidate(query_pattern=pat)
        elif op == "snapshot" and gpu._engine.shard is not None and not sharded:
            p = str(tmp_path / f"s{step}.json")
            gpu.save_snapshot(p)
            cpu._evict_expired()                        # save_snapshot drops expired entries first, like save()
            fresh = cache_mod.QueryCache(**kw)
            fresh.load_snapshot(p)
            assert [e.query for e in fresh._store.values()] == [e.query for e in gpu._store.values()]
            gpu = fresh                                 # carry on with the restored cache
        assert [e.query for e in gpu._store.values()] == [e.query for e in cpu._store.values()], (step, op)
        sa, sb = gpu.stats(), cpu.stats()
        assert (sa["size"], sa["valid"]) == (sb["size"], sb["valid"]), (step, op)
    assert hits > 20
