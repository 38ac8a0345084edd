"""GPU parity of the C-ABI hot path against the CPU oracle (seeded inputs, oracle-sized)."""
import numpy as np
import pytest

from oracle import qre_oracle as O
from qre_b200 import _cabi, synth

pytestmark = pytest.mark.gpu
REL = 1e-5          # north_star: scores within 1e-5 relative


def _load(w, max_batch=None, n_rows=None):
    n = w.rows.shape[0] if n_rows is None else n_rows
    sh = _cabi.Shard(dim=w.rows.shape[1], capacity=n + 64, max_batch=max_batch or w.queries.shape[0])
    slots = sh.append(w.rows[:n], w.row_ctx[:n], w.row_ts_us[:n], w.row_lru[:n], w.row_pred_dev[:n],
                      w.row_pred_conf[:n])
    assert np.array_equal(slots, np.arange(n))
    sh.set_centroids(w.centroids, w.owner)
    sh.set_perf_health(w.perf_health)
    return sh


def _oracle_scan(w, thr=0.85, n_rows=None):
    n = w.rows.shape[0] if n_rows is None else n_rows
    img = O.CacheImage(w.rows[:n], w.row_ctx[:n], w.row_ts_us[:n], w.row_lru[:n])
    return O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, thr)


def _check_lookup(out, exp_row, exp_sim):
    got = out["hit_slot"]
    bad = np.nonzero(got != exp_row)[0]
    assert bad.size == 0, (bad[:10], got[bad[:10]], exp_row[bad[:10]])
    hit = exp_row >= 0
    np.testing.assert_allclose(out["hit_sim"][hit], exp_sim[hit], rtol=REL)
    assert np.all(out["hit_sim"][~hit] == 0)


@pytest.fixture(scope="module")
def small():
    return synth.make_workload(6000, 300, 384, n_centroids=8, seed=3, frac_near=0.08, frac_dup=0.03)


def test_exact_scan_matches_oracle(small):
    w = small
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us, scan_mode=_cabi.SCAN_EXACT)
    out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us)
    exp_row, exp_sim = _oracle_scan(w)
    _check_lookup(out, exp_row, exp_sim)
    assert (exp_row >= 0).sum() > 100 and (exp_row < 0).sum() > 100
    hit = exp_row >= 0
    assert np.array_equal(out["hit_pred_dev"][hit], w.row_pred_dev[exp_row[hit]])
    assert np.array_equal(out["hit_pred_conf"][hit], w.row_pred_conf[exp_row[hit]])
    zero = ~np.any(w.queries != 0, axis=1)
    assert np.all(out["flags"][zero] & _cabi.FLAG_ZERO_NORM) and np.all(out["hit_slot"][zero] == -1)
    sh.close()


def test_tensor_gemm_raw_products(small):
    """tcgen05/TMA descriptor check: raw bf16 products against a CPU bf16 emulation."""
    import torch
    w = small
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us)
    m, n = w.queries.shape[0], 1500
    for row0 in (0, 4400):
        got = sh.debug_gemm(m, row0, n)
        qn = torch.from_numpy(w.queries / np.maximum(np.linalg.norm(w.queries, axis=1, keepdims=True), 1e-30))
        en = torch.from_numpy(w.rows[row0:row0 + n] /
                              np.maximum(np.linalg.norm(w.rows[row0:row0 + n], axis=1, keepdims=True), 1e-30))
        ref = (qn.bfloat16().float() @ en.bfloat16().float().T).numpy()
        # 1-ulp differences in the fp32 normalisation can flip a bf16 rounding: allow 2 bf16 ulps of slack
        np.testing.assert_allclose(got, ref, atol=3e-4, rtol=0)
        true = (qn.double() @ en.double().T).numpy()
        assert np.abs(got - true).max() < 2e-3        # well inside the eps the candidate window assumes
    sh.close()


def test_tensor_scan_matches_oracle(small):
    w = small
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us, want_cand=True)
    exp_row, exp_sim = _oracle_scan(w)
    _check_lookup(out, exp_row, exp_sim)
    cfg_e = _cabi.RouteConfig(ttl_us=w.ttl_us, scan_mode=_cabi.SCAN_EXACT)
    out_e = sh.lookup_route(w.queries, w.q_ctx, cfg_e, w.now_us)
    assert np.array_equal(out["hit_slot"], out_e["hit_slot"])
    assert np.array_equal(out["hit_sim"], out_e["hit_sim"])          # same summation order: bit-equal
    assert np.array_equal(out["cand"]["slot"][:, 0], out["hit_slot"])
    sh.close()


@pytest.mark.parametrize("m", [1, 7, 64, 129])
def test_tensor_scan_ragged_batches(small, m):
    w = small
    sh = _load(w, max_batch=256)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    out = sh.lookup_route(w.queries[:m], w.q_ctx[:m], cfg, w.now_us)
    exp_row, exp_sim = _oracle_scan(w)
    _check_lookup(out, exp_row[:m], exp_sim[:m])
    sh.close()


def test_route_matches_oracle(small):
    w = small
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us, tok=w.tok, heur_dev=w.heur_dev,
                          heur_conf=w.heur_conf)
    p = O.RouteParams()
    r = O.route_batch(w.queries, w.tok, w.heur_dev, w.heur_conf, w.centroids, w.owner, p)
    np.testing.assert_allclose(out["centroid_sim"], r["sims"], rtol=REL, atol=1e-7)
    assert np.array_equal(out["tok_dev"], r["tok_dev"]) and np.array_equal(out["tok_conf"], r["tok_conf"])
    # a decision may legitimately differ only where a margin sits within fp32 summation noise of a rule
    sims = np.sort(r["sims"].astype(np.float64).reshape(len(w.tok), -1, 1).max(axis=2), axis=1)
    dsim = np.full((len(w.tok), 2), -np.inf)
    for k in range(w.centroids.shape[0]):
        dsim[:, w.owner[k]] = np.maximum(dsim[:, w.owner[k]], r["sims"][:, k])
    margin = np.abs(dsim[:, 1] - dsim[:, 0])
    band = (np.abs(margin - 0.03) < 2e-6) | (np.abs(dsim.max(1) - 0.05) < 2e-6) | (margin < 2e-6)
    assert band.sum() <= 3
    ok = ~band
    assert np.array_equal(out["sem_method"][ok], r["sem_method"][ok])
    assert np.array_equal(out["sem_dev"][ok], r["sem_dev"][ok])
    assert np.array_equal(out["dev"][ok], r["dev"][ok])
    np.testing.assert_allclose(out["sem_conf"][ok], r["sem_conf"][ok], rtol=REL, atol=2e-6)
    np.testing.assert_allclose(out["conf"][ok], r["conf"][ok], rtol=REL, atol=4e-6)
    np.testing.assert_allclose(out["score"][ok], r["S"][ok], rtol=REL, atol=2e-6)
    assert set(np.unique(r["sem_method"])) == {0, 1, 2}
    del sims
    sh.close()


def test_update_tombstone_and_reuse(small):
    w = small
    n = 2000
    sh = _load(w, n_rows=n)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    exp_row, _ = _oracle_scan(w, n_rows=n)
    hits = np.nonzero(exp_row >= 0)[0]
    assert hits.size > 20
    victims = np.unique(exp_row[hits[:10]])
    sh.tombstone(victims)
    ctx2 = w.row_ctx[:n].copy(); ctx2[victims] = O.CTX_TOMBSTONE
    img = O.CacheImage(w.rows[:n], ctx2, w.row_ts_us[:n], w.row_lru[:n])
    exp2, sim2 = O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, 0.85)
    out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us)
    _check_lookup(out, exp2, sim2)
    # freed slots are reused (LIFO) and the new rows are found
    new_rows = w.rows[victims] * np.float32(0.5)
    slots = sh.append(new_rows, w.row_ctx[victims], w.row_ts_us[victims], w.row_lru[victims],
                      w.row_pred_dev[victims], w.row_pred_conf[victims])
    assert sorted(slots.tolist()) == sorted(victims.tolist())
    assert sh.info()["live"] == n
    rows3 = w.rows[:n].copy(); rows3[slots] = new_rows
    ctx3 = w.row_ctx[:n].copy(); ctx3[slots] = w.row_ctx[victims]
    ts3 = w.row_ts_us[:n].copy(); ts3[slots] = w.row_ts_us[victims]
    lru3 = w.row_lru[:n].copy(); lru3[slots] = w.row_lru[victims]
    img3 = O.CacheImage(rows3, ctx3, ts3, lru3)
    exp3, sim3 = O.scan_batch(w.queries, w.q_ctx, img3, w.now_us, w.ttl_us, 0.85)
    out3 = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us)
    _check_lookup(out3, exp3, sim3)
    # LRU promotion of one of two identical rows flips the winner (cache.py:292,303)
    np.testing.assert_array_equal(sh.read_rows(slots), new_rows)
    sh.close()


def test_lru_tie_break_and_ttl():
    d = 384
    rs = np.random.Generator(np.random.PCG64(9))
    base = rs.standard_normal((4, d)).astype(np.float32)
    rows = np.stack([base[0], base[0], base[1], base[2]])
    ctx = np.array([5, 5, 5, 6], np.uint64)
    ts = np.array([1000, 1000, 0, 1000], np.int64)
    lru = np.array([20, 10, 30, 40], np.uint64)
    sh = _cabi.Shard(dim=d, capacity=16, max_batch=8)
    sh.append(rows, ctx, ts, lru, np.zeros(4, np.uint8), np.ones(4))
    q = np.stack([base[0] * 3, base[1], base[2], base[2]])
    qc = np.array([5, 5, 6, 5], np.uint64)
    for mode in (_cabi.SCAN_TENSOR, _cabi.SCAN_EXACT):
        cfg = _cabi.RouteConfig(ttl_us=500, scan_mode=mode)
        out = sh.lookup_route(q, qc, cfg, now_us=1200)
        # twin rows: smaller lru_seq wins; row 2 expired (age 1200 > 500); ctx mismatch misses
        assert out["hit_slot"].tolist() == [1, -1, 3, -1], mode
        sh.update(np.array([1]), lru_seq=np.array([50], np.uint64))       # move_to_end
        out = sh.lookup_route(q, qc, cfg, now_us=1200)
        assert out["hit_slot"].tolist() == [0, -1, 3, -1], mode
        sh.update(np.array([1]), lru_seq=np.array([10], np.uint64))
        out = sh.lookup_route(q, qc, cfg, now_us=1500)                    # age == ttl is still valid
        assert out["hit_slot"].tolist() == [1, -1, 3, -1], mode
        out = sh.lookup_route(q, qc, cfg, now_us=1501)
        assert out["hit_slot"].tolist() == [-1, -1, -1, -1], mode
    sh.close()


def test_more_contenders_than_the_list_holds_are_chained():
    """More than QRE_TOPK near-identical contenders in one row range: the bf16 window cannot
    separate them; the ones that fall off the top-8 list are chained in spill chunks and re-scored
    exactly, so the answer equals the brute-force scan's (LRU-earliest of the best) without it."""
    d = 384
    rs = np.random.Generator(np.random.PCG64(10))
    e = rs.standard_normal(d).astype(np.float32)
    n = 40
    rows = np.stack([e * np.float32(1 + 0.01 * i) for i in range(n)])
    noise = rs.standard_normal((n, d)).astype(np.float32) * np.float32(1e-4)
    rows[5:] += noise[5:]
    ctx = np.full(n, 7, np.uint64)
    lru = np.arange(100, 100 + n, dtype=np.uint64)[::-1].copy()
    sh = _cabi.Shard(dim=d, capacity=64, max_batch=8)
    sh.append(rows, ctx, np.zeros(n, np.int64), lru, np.zeros(n, np.uint8), np.ones(n))
    q = e[None, :].copy()
    cfg = _cabi.RouteConfig(ttl_us=10**9)
    out = sh.lookup_route(q, np.array([7], np.uint64), cfg, now_us=1)
    img = O.CacheImage(rows, ctx, np.zeros(n, np.int64), lru)
    exp_row, exp_sim = O.scan_batch(q, np.array([7], np.uint64), img, 1, 10**9, 0.85)
    cfg_e = _cabi.RouteConfig(ttl_us=10**9, scan_mode=_cabi.SCAN_EXACT)
    out_e = sh.lookup_route(q, np.array([7], np.uint64), cfg_e, now_us=1)
    assert not (out["flags"][0] & _cabi.FLAG_OVERFLOW)
    # (the CPU oracle may name another of the rows whose scores tie within one ulp: OpenBLAS sums in
    # another order; the two GPU paths share one summation order and must agree exactly)
    assert out["hit_slot"][0] == out_e["hit_slot"][0]
    assert out["hit_sim"][0] == out_e["hit_sim"][0]
    np.testing.assert_allclose(out["hit_sim"][0], exp_sim[0], rtol=REL)
    sh.close()


def test_spill_pool_exhaustion_falls_back_to_exact_scan():
    """20,000 copies of one embedding against a handle whose spill pool holds 15,360 candidates
    (max_batch 8): the pool runs out, the query is flagged and the device-side brute-force pass
    settles it — still the exact answer (the LRU-earliest copy), on the asynchronous path too."""
    d, n = 384, 20000
    rs = np.random.Generator(np.random.PCG64(21))
    e = rs.standard_normal(d).astype(np.float32)
    rows = np.repeat(e[None, :], n, axis=0)
    ctx = np.full(n, 7, np.uint64)
    lru = rs.permutation(n).astype(np.uint64) + np.uint64(5)
    sh = _cabi.Shard(dim=d, capacity=n, max_batch=8)
    sh.append(rows, ctx, np.zeros(n, np.int64), lru, np.zeros(n, np.uint8), np.ones(n))
    q = np.stack([e, rs.standard_normal(d).astype(np.float32)])
    qc = np.array([7, 7], np.uint64)
    cfg = _cabi.RouteConfig(ttl_us=10**9)
    out = sh.lookup_route(q, qc, cfg, now_us=1)
    assert out["flags"][0] & _cabi.FLAG_OVERFLOW and not (out["flags"][1] & _cabi.FLAG_OVERFLOW)
    assert out["hit_slot"].tolist() == [int(np.argmin(lru)), -1]
    sh.lookup_route_begin(0, q, qc, cfg, 1)
    sh.lookup_route_begin(1, q[::-1].copy(), qc, cfg, 1)
    a, b = sh.lookup_route_end(0), sh.lookup_route_end(1)
    assert a["hit_slot"].tolist() == [int(np.argmin(lru)), -1] and b["hit_slot"].tolist() == [-1, int(np.argmin(lru))]
    sh.close()


def test_config4_768d_64_centroids_16_devices_perf_term():
    """BASELINE config 4 at 262,144 rows: 768-d rows (streaming-A scan variant), 64 centroids owned
    by 16 devices, perf health vector in the hybrid score."""
    w = synth.make_workload(262144, 260, 768, n_centroids=64, n_devices=16, seed=8, frac_near=0.06, frac_dup=0.03)
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us, n_devices=16, w_perf=0.1)
    out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us, tok=w.tok, heur_dev=w.heur_dev,
                          heur_conf=w.heur_conf)
    exp_row, exp_sim = _oracle_scan(w)
    _check_lookup(out, exp_row, exp_sim)
    assert (exp_row >= 0).sum() > 80
    got = sh.debug_gemm(260, 256, 512)
    qn = w.queries / np.maximum(np.linalg.norm(w.queries, axis=1, keepdims=True), 1e-30)
    en = w.rows[256:768] / np.maximum(np.linalg.norm(w.rows[256:768], axis=1, keepdims=True), 1e-30)
    assert np.abs(got - qn.astype(np.float64) @ en.astype(np.float64).T).max() < 2e-3
    p = O.RouteParams(n_devices=16, w_perf=0.1)
    r = O.route_batch(w.queries, w.tok, w.heur_dev, w.heur_conf, w.centroids, w.owner, p, w.perf_health)
    np.testing.assert_allclose(out["centroid_sim"], r["sims"], rtol=REL, atol=1e-7)
    dsim = np.full((260, 16), -np.inf)
    for k in range(64):
        dsim[:, w.owner[k]] = np.maximum(dsim[:, w.owner[k]], r["sims"][:, k])
    top2 = np.sort(dsim, axis=1)[:, -2:]
    margin = top2[:, 1] - top2[:, 0]
    S2 = np.sort(r["S"], axis=1)[:, -2:]
    band = (np.abs(margin - 0.03) < 2e-6) | (np.abs(top2[:, 1] - 0.05) < 2e-6) | (margin < 2e-6) \
        | (np.abs(S2[:, 1] - S2[:, 0]) < 1e-6)
    ok = ~band
    assert band.sum() <= 4
    assert np.array_equal(out["sem_method"][ok], r["sem_method"][ok])
    assert np.array_equal(out["sem_dev"][ok], r["sem_dev"][ok])
    assert np.array_equal(out["dev"][ok], r["dev"][ok])
    np.testing.assert_allclose(out["score"][ok], r["S"][ok], rtol=REL, atol=2e-6)
    np.testing.assert_allclose(out["conf"][ok], r["conf"][ok], rtol=REL, atol=4e-6)
    assert len(np.unique(r["dev"])) > 3
    sh.close()


def test_pipelined_slots_single_gpu(small):
    """Double-buffered query state: the merge of batch i on a side stream while the scan of batch
    i+1 runs must give the serial results (world = 1: no collective, same code path otherwise)."""
    import torch
    from qre_b200.sharded import ShardedLookup
    w = small
    sh = _load(w)
    dev = torch.device("cuda", 0)
    look = ShardedLookup(sh, _cabi.RouteConfig(ttl_us=w.ttl_us), 0, 1, reserve_sms=8)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    perms = [np.random.Generator(np.random.PCG64(s_)).permutation(300) for s_ in range(4)]
    ins = [(t(w.queries[pm]), t(w.q_ctx[pm].view(np.int64)), t(w.tok[pm]), t(w.heur_dev[pm]), t(w.heur_conf[pm]))
           for pm in perms]
    serial = []
    for a, b, c, d, e in ins:
        o = look.step(a, b, w.now_us, c, d, e)
        torch.cuda.synchronize()
        serial.append({k: v.cpu().numpy().copy() for k, v in o.items()})
    exp_row, _ = _oracle_scan(w)
    assert np.array_equal(serial[0]["hit_slot"], exp_row[perms[0]])
    torch.cuda.synchronize()
    got = []
    for i, (a, b, c, d, e) in enumerate(ins):
        o, ev = look.step_pipelined(a, b, w.now_us, c, d, e)
        if i >= 1:      # consume batch i-1 before its buffers are reused by batch i+1
            po, pev = prev
            pev.synchronize()
            got.append({k: v.cpu().numpy().copy() for k, v in po.items()})
        prev = (o, ev)
    prev[1].synchronize()
    got.append({k: v.cpu().numpy().copy() for k, v in prev[0].items()})
    for i in range(4):
        for k in ("hit_slot", "hit_sim", "dev", "conf", "score", "sem_method", "flags"):
            assert np.array_equal(got[i][k], serial[i][k]), (i, k)
    sh.close()


def test_many_queries_with_crowded_ranges():
    """20 queries, each facing 12 near-identical rows (more than a list holds): chained candidates,
    no brute-force pass; the async device pipeline must return the exact-scan answers on its own."""
    import torch
    from qre_b200.sharded import ShardedLookup
    d, nq, per = 384, 20, 12
    rs = np.random.Generator(np.random.PCG64(12))
    base = rs.standard_normal((nq, d)).astype(np.float32)
    rows = np.repeat(base, per, axis=0) + rs.standard_normal((nq * per, d)).astype(np.float32) * np.float32(2e-4)
    filler = rs.standard_normal((3000, d)).astype(np.float32)
    rows = np.concatenate([rows, filler])
    n = rows.shape[0]
    ctx = np.full(n, 3, np.uint64)
    lru = rs.permutation(n).astype(np.uint64) + np.uint64(1)
    sh = _cabi.Shard(dim=d, capacity=n, max_batch=32)
    sh.append(rows, ctx, np.zeros(n, np.int64), lru, np.zeros(n, np.uint8), np.ones(n))
    q = np.concatenate([base, rs.standard_normal((4, d)).astype(np.float32)])       # 4 plain misses
    qc = np.full(q.shape[0], 3, np.uint64)
    out_e = sh.lookup_route(q, qc, _cabi.RouteConfig(ttl_us=10**9, scan_mode=_cabi.SCAN_EXACT), now_us=1)
    out_t = sh.lookup_route(q, qc, _cabi.RouteConfig(ttl_us=10**9), now_us=1)
    assert not np.any(out_t["flags"] & _cabi.FLAG_OVERFLOW)
    assert np.array_equal(out_t["hit_slot"], out_e["hit_slot"]) and np.array_equal(out_t["hit_sim"], out_e["hit_sim"])
    assert np.all(out_e["hit_slot"][:nq] >= 0) and np.all(out_e["hit_slot"][nq:] == -1)
    dev = torch.device("cuda", 0)
    look = ShardedLookup(sh, _cabi.RouteConfig(ttl_us=10**9), 0, 1)
    o = look.step(torch.from_numpy(q).to(dev), torch.from_numpy(qc.view(np.int64)).to(dev), 1)
    torch.cuda.synchronize()
    assert np.array_equal(o["hit_slot"].cpu().numpy(), out_e["hit_slot"])
    assert np.array_equal(o["hit_sim"].cpu().numpy(), out_e["hit_sim"])
    assert int(look.ovf.item()) == 0            # nothing left for the host to settle
    sh.close()


def test_begin_end_slots_overlap_and_pinned_inputs(small):
    """Two batches in flight on the two host slots, one from page-locked and one from pageable
    memory: each equals the synchronous call."""
    import torch
    w = small
    sh = _load(w)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    parts = [slice(0, 150), slice(150, 300)]
    sync = [sh.lookup_route(w.queries[p], w.q_ctx[p], cfg, w.now_us, tok=w.tok[p], heur_dev=w.heur_dev[p],
                            heur_conf=w.heur_conf[p]) for p in parts]
    pinned_q = torch.from_numpy(np.ascontiguousarray(w.queries[parts[0]])).pin_memory().numpy()
    sh.lookup_route_begin(0, pinned_q, w.q_ctx[parts[0]], cfg, w.now_us, tok=w.tok[parts[0]],
                          heur_dev=w.heur_dev[parts[0]], heur_conf=w.heur_conf[parts[0]])
    sh.lookup_route_begin(1, w.queries[parts[1]], w.q_ctx[parts[1]], cfg, w.now_us, tok=w.tok[parts[1]],
                          heur_dev=w.heur_dev[parts[1]], heur_conf=w.heur_conf[parts[1]])
    with pytest.raises(_cabi.QreError, match="still has a batch in flight"):
        sh.lookup_route_begin(1, w.queries[parts[1]], w.q_ctx[parts[1]], cfg, w.now_us)
    got = [sh.lookup_route_end(0), sh.lookup_route_end(1)]
    for g, s_ in zip(got, sync):
        for k in s_:
            assert np.array_equal(g[k], s_[k]), k
    import ctypes
    assert sh.lib.qre_lookup_route_end(sh.h, 0, ctypes.byref(_cabi.Result())) == 1
    assert b"no batch in flight" in sh.lib.qre_last_error()
    sh.close()


def test_three_batches_in_flight_and_a_stream_of_them(small):
    """Three host slots: three batches in flight at once, then a run of 9 batches kept two ahead
    (begin i+2 before end i, the multi-rank calling pattern); every result equals the synchronous
    call.  The large batches alternate between the handle's two scan streams."""
    w = small
    big = synth.make_workload(6000, 520, 384, n_centroids=8, seed=21)          # > one query block: pipelined path
    for wk, step in ((w, 100), (big, 260)):
        sh = _load(wk)
        cfg = _cabi.RouteConfig(ttl_us=wk.ttl_us)
        parts = [slice(i * step // 2, i * step // 2 + step) for i in range(3)]
        sync = [sh.lookup_route(wk.queries[p], wk.q_ctx[p], cfg, wk.now_us, tok=wk.tok[p], heur_dev=wk.heur_dev[p],
                                heur_conf=wk.heur_conf[p]) for p in parts]

        def begin(slot, p):
            sh.lookup_route_begin(slot, wk.queries[p], wk.q_ctx[p], cfg, wk.now_us, tok=wk.tok[p],
                                  heur_dev=wk.heur_dev[p], heur_conf=wk.heur_conf[p])
        for i, p in enumerate(parts):
            begin(i, p)
        with pytest.raises(_cabi.QreError, match="slot must be 0, 1 or 2"):
            begin(3, parts[0])
        for i in (2, 0, 1):                              # any order of completion
            g = sh.lookup_route_end(i)
            for k in sync[i]:
                assert np.array_equal(g[k], sync[i][k]), (i, k)
        n = 9
        for i in range(n + 2):
            if i < n:
                begin(i % 3, parts[i % 3])
            if i >= 2:
                g = sh.lookup_route_end((i - 2) % 3)
                for k in sync[(i - 2) % 3]:
                    assert np.array_equal(g[k], sync[(i - 2) % 3][k]), (i, k)
        sh.close()


def test_group_lookup_route_two_shards_one_process(small):
    """qre_group_lookup_route: one process, the rows dealt over two handles (two GPUs when the box has
    them, else both on GPU 0), lists gathered and merged on the first handle's GPU.  Must equal the
    oracle over the whole cache, including the cross-shard tie-break by LRU order."""
    import torch
    w = small
    devs = [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]
    n = w.rows.shape[0]
    rows = w.rows.copy()
    ctx = w.row_ctx.copy()
    rows[1:41:2] = rows[0:40:2]               # twins in different shards: lru decides
    ctx[1:41:2] = ctx[0:40:2]
    shards = []
    for r, dev in enumerate(devs):
        idx = np.arange(r, n, 2)
        sh = _cabi.Shard(dim=384, capacity=idx.size, max_batch=300, device=dev)
        sh.append(rows[idx], ctx[idx], w.row_ts_us[idx], w.row_lru[idx], w.row_pred_dev[idx], w.row_pred_conf[idx])
        sh.set_centroids(w.centroids, w.owner)
        shards.append(sh)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    out = _cabi.group_lookup_route(shards, w.queries, w.q_ctx, cfg, w.now_us, tok=w.tok, heur_dev=w.heur_dev,
                                   heur_conf=w.heur_conf)
    img = O.CacheImage(rows, ctx, w.row_ts_us, w.row_lru)
    exp_row, exp_sim = O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, 0.85)
    glob = np.where(out["hit_slot"] >= 0, out["hit_slot"] * 2 + out["hit_shard"], -1)
    assert np.array_equal(glob, exp_row)
    hit = exp_row >= 0
    assert hit.sum() > 50
    np.testing.assert_allclose(out["hit_sim"][hit], exp_sim[hit], rtol=REL)
    assert np.array_equal(out["hit_pred_dev"][hit], w.row_pred_dev[exp_row[hit]])
    r = O.route_batch(w.queries, w.tok, w.heur_dev, w.heur_conf, w.centroids, w.owner, O.RouteParams())
    np.testing.assert_allclose(out["centroid_sim"], r["sims"], rtol=REL, atol=1e-7)
    assert float(np.mean(out["dev"] == r["dev"])) > 0.995
    out2 = _cabi.group_lookup_route(shards, w.queries, w.q_ctx, cfg, w.now_us)      # lookup only, again
    assert np.array_equal(out2["hit_slot"], out["hit_slot"]) and np.array_equal(out2["hit_sim"], out["hit_sim"])
    for sh in shards:
        sh.close()
