"""GPU: qre_exact_lookup_noalloc against the existing entry point (qre_lookup_route) and the oracle.

The no-allocation entry point exists for hosts that may not call cudaMalloc when a lookup
arrives (all HBM already handed to their own pool, request threads that cannot absorb the
device-wide stall of an allocation, plugins confined to a host framework's allocator).  It
takes every array from the caller and keeps its scratch in a static device array inside the
library, so this test owns all device memory itself (torch tensors stand in for the caller's
pool) and only checks results: the hit row and score must equal what the handle-based entry
point returns for the same cache and queries, and what the CPU oracle computes, to the tolerance
the other parity tests use (identical decisions, scores within 1e-5 relative).

Two calls in a row with different queries: the second call sees the scratch exactly as the
first one left it, which is the situation the "clear what must start at zero" rule is about.

Collected last on purpose (file name), so a run that stops at the first failure has already
run every other test; it needs nothing from them and also passes alone in a fresh process.
"""
import numpy as np
import pytest

from oracle import qre_oracle as O
from qre_b200 import _cabi, synth

pytestmark = pytest.mark.gpu
REL = 1e-5


def _existing(w, queries, q_ctx):
    sh = _cabi.Shard(dim=w.rows.shape[1], capacity=w.rows.shape[0], max_batch=8)
    sh.append(w.rows, w.row_ctx, w.row_ts_us, w.row_lru, w.row_pred_dev, w.row_pred_conf)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    out = sh.lookup_route(queries, q_ctx, cfg, w.now_us)
    sh.close()
    return out["hit_slot"], out["hit_sim"]


@pytest.mark.parametrize("dim,m,own_stream", [(384, 8, False), (1024, 5, True), (384, 1, True)])
def test_noalloc_matches_existing_entry_point_twice(dim, m, own_stream):
    import torch
    dev = torch.device("cuda", 0)
    w = synth.make_workload(20000 if dim == 384 else 6000, 16, dim, n_centroids=2, seed=17 + dim + m,
                            frac_planted=0.6, frac_near=0.1, frac_dup=0.1, n_zero_q=1, n_zero_rows=2)
    img = O.CacheImage(w.rows, w.row_ctx, w.row_ts_us, w.row_lru)
    # the caller's memory: everything the entry point reads or writes
    rows = torch.from_numpy(w.rows).to(dev)
    ctx = torch.from_numpy(w.row_ctx.view(np.int64)).to(dev)
    ts = torch.from_numpy(w.row_ts_us).to(dev)
    lru = torch.from_numpy(w.row_lru.view(np.int64)).to(dev)
    hit_row = torch.full((8,), -7, dtype=torch.int64, device=dev)
    hit_sim = torch.full((8,), -7.0, dtype=torch.float32, device=dev)
    st = torch.cuda.Stream(device=dev) if own_stream else None
    n_hits = 0
    for call in range(2):                                   # two calls in a row, different queries
        q_np = np.ascontiguousarray(w.queries[call * 8:call * 8 + m])
        c_np = np.ascontiguousarray(w.q_ctx[call * 8:call * 8 + m])
        q = torch.from_numpy(q_np).to(dev)
        qc = torch.from_numpy(c_np.view(np.int64)).to(dev)
        torch.cuda.synchronize()
        _cabi.exact_lookup_noalloc(0, rows.shape[0], dim, rows.data_ptr(), ctx.data_ptr(), ts.data_ptr(),
                                   lru.data_ptr(), m, q.data_ptr(), qc.data_ptr(), 0.85, w.now_us, w.ttl_us,
                                   hit_row.data_ptr(), hit_sim.data_ptr(),
                                   stream=st.cuda_stream if st is not None else 0)
        torch.cuda.synchronize()
        exp_row, exp_sim = _existing(w, q_np, c_np)
        got_row, got_sim = hit_row.cpu().numpy()[:m], hit_sim.cpu().numpy()[:m]
        assert np.array_equal(got_row, exp_row), (call, got_row, exp_row)
        np.testing.assert_allclose(got_sim, exp_sim, rtol=REL)
        # and the CPU restatement of src/cache.py:280-296 directly
        o_row, o_sim = O.scan_batch(q_np, c_np, img, w.now_us, w.ttl_us, 0.85)
        assert np.array_equal(got_row, o_row), (call, got_row, o_row)
        np.testing.assert_allclose(got_sim[o_row >= 0], o_sim[o_row >= 0], rtol=REL)
        n_hits += int((exp_row >= 0).sum())
    assert 1 <= n_hits <= 2 * m - (1 if m > 1 else 0)       # hits were exercised (and misses when m > 1)


def test_noalloc_rejects_sizes_beyond_the_maximum():
    import torch
    dev = torch.device("cuda", 0)
    z = torch.zeros(4096, dtype=torch.float32, device=dev)
    zi = torch.zeros(16, dtype=torch.int64, device=dev)
    args = dict(device=0, n_rows=1, dim=384, rows_ptr=z.data_ptr(), row_ctx_ptr=zi.data_ptr(),
                row_ts_ptr=zi.data_ptr(), row_lru_ptr=zi.data_ptr(), m=8, q_ptr=z.data_ptr(),
                q_ctx_ptr=zi.data_ptr(), threshold=0.85, now_us=0, ttl_us=1, hit_row_ptr=zi.data_ptr(),
                hit_sim_ptr=z.data_ptr())
    with pytest.raises(_cabi.QreError, match=r"m must be in \[1, 8\]"):
        _cabi.exact_lookup_noalloc(**{**args, "m": 9})
    with pytest.raises(_cabi.QreError, match="dim must be a multiple of 4"):
        _cabi.exact_lookup_noalloc(**{**args, "dim": 1028})
    with pytest.raises(_cabi.QreError, match="n_rows"):
        _cabi.exact_lookup_noalloc(**{**args, "n_rows": 2**31})
    # a slice of a pool allocation that is not 16-byte aligned: an error code, not a device fault
    with pytest.raises(_cabi.QreError, match="16-byte aligned"):
        _cabi.exact_lookup_noalloc(**{**args, "rows_ptr": z.data_ptr() + 4})
    torch.cuda.synchronize()
