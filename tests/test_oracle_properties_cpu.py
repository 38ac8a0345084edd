"""CPU property tests (hypothesis) of the oracle itself: the vectorised scan used as the fast
checker at 1M rows must agree with the faithful per-entry loop on arbitrary small instances —
including duplicate rows (LRU tie-break), expired and tombstoned rows, other-context rows, zero
vectors and thresholds on either side of zero — and obey the metric's invariances."""
import numpy as np
from hypothesis import given, settings, strategies as st

from oracle import qre_oracle as O


@st.composite
def instances(draw):
    seed = draw(st.integers(0, 2**31 - 1))
    n = draw(st.integers(1, 40))
    m = draw(st.integers(1, 6))
    d = draw(st.sampled_from([8, 16, 64]))
    thr = draw(st.sampled_from([0.85, 0.5, 0.0, -0.3]))
    rs = np.random.Generator(np.random.PCG64(seed))
    rows = rs.standard_normal((n, d)).astype(np.float32)
    for _ in range(draw(st.integers(0, 4))):                      # duplicates and scaled copies
        i, j = rs.integers(0, n, 2)
        rows[j] = rows[i] * np.float32(rs.choice([1.0, 2.0, 0.5]))
    if draw(st.booleans()):
        rows[rs.integers(0, n)] = 0.0
    ctx = rs.integers(1, 3, n).astype(np.uint64)
    ctx[rs.random(n) < 0.1] = O.CTX_TOMBSTONE
    ts = rs.integers(0, 100, n).astype(np.int64)
    lru = (rs.permutation(n) + 1).astype(np.uint64)
    Q = rs.standard_normal((m, d)).astype(np.float32)
    for q in range(m):
        if rs.random() < 0.6:
            Q[q] = rows[rs.integers(0, n)] * np.float32(rs.uniform(0.5, 2)) \
                + np.float32(rs.choice([0.0, 0.05])) * rs.standard_normal(d).astype(np.float32)
    if draw(st.booleans()):
        Q[0] = 0.0
    qctx = rs.integers(1, 3, m).astype(np.uint64)
    return O.CacheImage(rows, ctx, ts, lru), Q, qctx, thr, int(rs.integers(50, 150)), int(rs.integers(10, 120))


@settings(max_examples=150, deadline=None)
@given(instances())
def test_batch_scan_equals_faithful_scan(inst):
    img, Q, qctx, thr, now, ttl = inst
    rows_b, sims_b = O.scan_batch(Q, qctx, img, now, ttl, thr)
    for q in range(Q.shape[0]):
        r, s = O.scan_faithful(Q[q], int(qctx[q]), img, now, ttl, thr)
        if r != rows_b[q]:
            # only an fp32 summation-order tie between two rows may flip the winner
            assert r >= 0 and rows_b[q] >= 0
            assert abs(O.cache_similarity(Q[q], img.rows[r]) - O.cache_similarity(Q[q], img.rows[rows_b[q]])) < 2e-6
        else:
            assert abs(float(sims_b[q]) - s) <= 2e-6 * max(1.0, abs(s))
        if r >= 0:
            assert s > 0 and s >= thr and img.ctx[r] == qctx[q] and now - img.ts_us[r] <= ttl


@settings(max_examples=60, deadline=None)
@given(instances(), st.floats(0.25, 4.0))
def test_scale_invariance_and_topk_prefix(inst, scale):
    img, Q, qctx, thr, now, ttl = inst
    r1, s1 = O.scan_batch(Q, qctx, img, now, ttl, thr)
    r2, s2 = O.scan_batch(Q * np.float32(scale), qctx, img, now, ttl, thr)
    same = r1 == r2
    assert same.mean() >= 0.8                          # cosine ignores the query's scale (up to fp32 ties)
    np.testing.assert_allclose(s1[same], s2[same], rtol=5e-6, atol=1e-7)
    r3, s3, top_r, top_s = O.scan_batch(Q, qctx, img, now, ttl, thr, topk=4)
    assert np.array_equal(r3, top_r[:, 0]) and np.array_equal(r1, r3)
    for q in range(Q.shape[0]):                         # valid entries form a prefix, sorted best first
        k = int((top_r[q] >= 0).sum())
        assert np.all(top_r[q, :k] >= 0) and np.all(top_r[q, k:] == -1)
        assert np.all(np.diff(top_s[q, :k].astype(np.float64)) <= 0)
