"""CPU: the batch forms of the host bookkeeping equal their per-entry forms.

* `QueryCache.insert_batch` against a loop of `QueryCache.insert` (reference cache.py:328-370): same
  store order, same entries, same device mirror, through duplicates inside the batch, refreshes of
  existing keys and evictions at capacity.
* `qre_host_features_packed` against `qre_host_features`, and its exact-match keys against
  `QueryCache._make_hash` (reference cache.py:518-521: md5 of "<ctx>||<query lowered, stripped>")."""
import hashlib

import numpy as np
import pytest

from oracle import ref_harness
from tests import replay_common as R
from tests.fake_engine import OracleEngine


def _mirror(qc):
    """The device rows as the host sees them: key -> (ctx, ts, lru, pdev, pconf, row)."""
    eng = qc._engine
    names = {v: k for k, v in qc._ctx_ids.items()}               # the ids themselves are arbitrary
    out = {}
    for slot, key in qc._slot_key.items():
        out[key] = (names[int(eng.ctx[slot])], int(eng.ts[slot]), int(eng.lru[slot]), int(eng.pdev[slot]),
                    float(eng.pconf[slot]), eng.rows[slot].copy())
    return out


def _same_state(a, b):
    assert list(a._store) == list(b._store)                      # LRU order
    for k in a._store:
        ea, eb = a._store[k], b._store[k]
        assert (ea.query, ea.context_key, ea.device_used, ea.hit_count, ea.response_time, ea._seq) == \
               (eb.query, eb.context_key, eb.device_used, eb.hit_count, eb.response_time, eb._seq), k
        assert ea.timestamp == eb.timestamp
        assert [r.to_dict() for r in ea.routing_history] == [r.to_dict() for r in eb.routing_history]
        assert ea.predict_device() == eb.predict_device()
        assert np.array_equal(ea.embedding, eb.embedding)
        assert ea.to_dict() == eb.to_dict()
    ma, mb = _mirror(a), _mirror(b)
    assert set(ma) == set(mb) == set(a._store)
    for k in ma:
        assert ma[k][:5] == mb[k][:5], k
        assert np.array_equal(ma[k][5], mb[k][5])
    assert a.stats() == b.stats()


@pytest.mark.parametrize("max_size", [64, 12])
def test_insert_batch_equals_insert_loop(max_size):
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(7))
    d = 32
    a = cache_mod.QueryCache(max_size=max_size, ttl_seconds=50, engine=OracleEngine(capacity=max_size, dim=d))
    b = cache_mod.QueryCache(max_size=max_size, ttl_seconds=50, engine=OracleEngine(capacity=max_size, dim=d))
    for rnd in range(6):
        n = 10
        qs = [f"q{rs.integers(0, 30)}" for _ in range(n)]        # repeats inside and across batches
        if rnd == 2:
            qs[3] = qs[1]
        cks = [("c0" if rs.random() < 0.7 else "c1") for _ in range(n)]
        devs = [("orin" if rs.random() < 0.5 else "nano") for _ in range(n)]
        confs = [float(rs.random()) for _ in range(n)]
        if rnd == 1:
            confs[0] = 0.0                                       # weightless record: prediction confidence 0.5
        emb = rs.standard_normal((n, d)).astype(np.float32)
        for i in range(n):
            a.insert(qs[i], cks[i], devs[i], confs[i], "hybrid", emb[i])
        b.insert_batch(qs, cks, devs, confs, "hybrid", emb)
        _same_state(a, b)
        clock.advance(7)
        # lookups (exact and semantic) see the same thing afterwards
        for i in (0, 4, 9):
            ra, rb = a.lookup(qs[i], cks[i], emb[i]), b.lookup(qs[i], cks[i], emb[i])
            assert (ra is None) == (rb is None)
            if ra is not None:
                assert (ra.entry.query, ra.predicted_device, ra.predicted_confidence) == \
                       (rb.entry.query, rb.predicted_device, rb.predicted_confidence)
        clock.advance(9)                                         # some entries expire over the rounds
    emb[0, 0] = 123.0                                            # the cache keeps its own copy of the block
    assert all(e.embedding[0] != 123.0 for e in b._store.values())


def test_insert_batch_expiry_records_cover_every_entry():
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(8))
    qc = cache_mod.QueryCache(max_size=100, ttl_seconds=10, engine=OracleEngine(capacity=100, dim=16))
    emb = rs.standard_normal((20, 16)).astype(np.float32)
    qc.insert_batch([f"a{i}" for i in range(20)], ["c"] * 20, ["nano"] * 20, [0.9] * 20, "hybrid", emb)
    clock.advance(5)
    qc.insert("fresh", "c", "orin", 0.9, "hybrid", emb[0])
    qc.insert("a3", "c", "orin", 0.9, "hybrid", emb[3])          # refreshed: must survive the sweep below
    clock.advance(6)
    assert qc.lookup("nothing", "c", emb[1]) is None             # sweeps the 19 expired entries
    assert sorted(qc._store) == sorted(qc._make_hash(q, "c") for q in ("fresh", "a3"))
    assert qc.stats()["evictions"] == 19


def test_packed_features_equal_pointer_form_and_keys_equal_md5():
    from qre_b200 import _cabi
    rs = np.random.Generator(np.random.PCG64(9))
    words = ["implement", "What is", " hello ", "compare", "daily intake", "wéather", "Traceback", "2 + 2 ?",
             "step by step", "\tdefine  ", "SELECT ;", "vs", "", "ÅNGSTRÖM", "ΣΟΦΙΑ"]
    qs = []
    for i in range(3000):
        k = int(rs.integers(1, 6))
        q = " ".join(words[int(j)] for j in rs.integers(0, len(words), k)) + ("x" * int(rs.integers(0, 140)))
        qs.append(q.upper() if i % 7 == 0 else q)
    cks = [("default", "ctx-ü", "")[int(j)] for j in rs.integers(0, 3, len(qs))]
    clen = rs.integers(0, 2000, len(qs)).tolist()
    hp = _cabi.HeurParams(240, 3, 2, 800, 1000, 2)
    a = _cabi.host_features(qs, clen, hp)
    b = _cabi.host_features_packed(qs, clen, hp, context_keys=cks, n_threads=3)
    for k in ("tok", "heur_dev", "heur_conf", "rule", "cat"):
        assert np.array_equal(a[k], b[k]), k
    n_none = 0
    for q, c, key, rule in zip(qs, cks, b["keys"], b["rule"]):
        assert rule >= 0 and key == hashlib.md5(f"{c}||{q.lower().strip()}".encode("utf-8")).hexdigest(), repr(q)
    assert n_none == 0
    # without context keys no key is computed; an empty batch is fine
    assert "keys" not in _cabi.host_features_packed(qs[:5], clen[:5], hp)
    assert _cabi.host_features_packed([], [], hp, context_keys=[])["keys"] == []


def test_warm_up_equals_insert_loop_with_and_without_vectors():
    """warm_up (reference cache.py:398-420) goes through insert_batch when the embedder delivers; an
    embedder that fails leaves entries without vectors, as in the reference."""
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(12))
    pairs = [(f"warm {i % 9}", "c" if i % 3 else "d", "orin" if i % 2 else "nano") for i in range(14)]   # repeats refresh
    table = {q: rs.standard_normal(24).astype(np.float32) for q, _, _ in pairs}

    class Emb:
        def encode(self, texts):
            return np.stack([table[t] for t in texts])

    class Broken:
        def encode(self, texts):
            raise RuntimeError("no model")

    a = cache_mod.QueryCache(max_size=32, ttl_seconds=50, engine=OracleEngine(capacity=32, dim=24))
    b = cache_mod.QueryCache(max_size=32, ttl_seconds=50, engine=OracleEngine(capacity=32, dim=24))
    a.warm_up(pairs, Emb())
    for q, c, d in pairs:
        b.insert(q, c, d, q_emb=table[q])
    _same_state(a, b)
    r = a.lookup("warm 4 paraphrase", "c", table["warm 4"] * 3)
    assert r is not None and r.entry.query == "warm 4"
    e = cache_mod.QueryCache(max_size=32, ttl_seconds=50, engine=OracleEngine(capacity=32, dim=24))
    e.warm_up(pairs, Broken())
    assert len(e._store) == len({(q, c) for q, c, _ in pairs}) and all(x.embedding is None for x in e._store.values())
    assert e.lookup("warm 4", "c") is not None and e.lookup("warm 4 paraphrase", "c", table["warm 4"]) is None


def test_load_replaces_rows_in_one_deferred_flush(tmp_path):
    """load() (reference cache.py:434-465) over a cache that already holds some of the keys: the
    file's rows replace them (the old vectors no longer hit), new keys are appended, entries that
    are not in the file stay, and a file larger than the store grows it — all mirrored to the device
    through one deferred flush."""
    clock = ref_harness.FakeClock()
    cache_mod, _, _ = R.load_modules(clock, OracleEngine)
    rs = np.random.Generator(np.random.PCG64(13))
    d = 16
    va, vb = rs.standard_normal((6, d)).astype(np.float32), rs.standard_normal((4, d)).astype(np.float32)
    a = cache_mod.QueryCache(max_size=8, ttl_seconds=100, engine=OracleEngine(capacity=8, dim=d))
    for i in range(6):
        a.insert(f"k{i}", "c", "orin", 0.9, "hybrid", va[i])
    path = str(tmp_path / "a.json")
    a.save(path)
    b = cache_mod.QueryCache(max_size=4, ttl_seconds=100, engine=OracleEngine(capacity=4, dim=d))
    for i, q in enumerate(["k1", "k4", "own0", "own1"]):
        b.insert(q, "c", "nano", 0.9, "hybrid", vb[i])
    assert b.load(path) == 6 and len(b._store) == 8               # 4 + 6 - 2 shared keys; grew past max_size
    for i in range(6):                                            # every file row is what the device now scans
        r = b.lookup("paraphrase", "c", va[i] * 2)
        assert r is not None and r.entry.query == f"k{i}" and r.predicted_device == "orin"
    for i, q in enumerate(["k1", "k4"]):                          # the replaced vectors are gone
        r = b.lookup("paraphrase", "c", vb[i])
        assert r is None or r.entry.query != q
    for i, q in enumerate(["own0", "own1"], start=2):             # untouched entries still hit
        r = b.lookup("paraphrase", "c", vb[i])
        assert r is not None and r.entry.query == q
    assert sorted(b._slot_key.values()) == sorted(b._store)
