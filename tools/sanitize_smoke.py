"""Small end-to-end run for compute-sanitizer: every kernel of the library on tiny shapes
(single-CTA scan, CTA-pair scan, streaming-A scan, re-score, overflow settlement, exact scan,
merge+route, row maintenance, no-alloc lookup)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]
from qre_b200 import _cabi, synth  # noqa: E402


def run(dim, n, m, k, d_):
    w = synth.make_workload(n, m, dim, n_centroids=k, n_devices=d_, seed=dim + m, frac_dup=0.05)
    sh = _cabi.Shard(dim=dim, capacity=n + 8, max_batch=m)
    sh.append(w.rows, w.row_ctx, w.row_ts_us, w.row_lru, w.row_pred_dev, w.row_pred_conf)
    sh.set_centroids(w.centroids, w.owner)
    sh.set_perf_health(w.perf_health)
    for mode in (_cabi.SCAN_TENSOR, _cabi.SCAN_EXACT):
        cfg = _cabi.RouteConfig(ttl_us=w.ttl_us, n_devices=d_, scan_mode=mode, w_perf=0.1)
        out = sh.lookup_route(w.queries, w.q_ctx, cfg, w.now_us, tok=w.tok, heur_dev=w.heur_dev,
                              heur_conf=w.heur_conf, want_cand=True)
    sh.tombstone(np.arange(0, 16))
    sh.update(np.array([20, 21]), rows=w.rows[:2], lru_seq=np.array([9, 8], np.uint64))
    sh.read_rows(np.array([20, 21, 22]))
    sh.lookup_route(w.queries[:1], w.q_ctx[:1], _cabi.RouteConfig(ttl_us=w.ttl_us, similarity_threshold=0.0), w.now_us)
    sh.close()
    return int((out["hit_slot"] >= 0).sum())


if __name__ == "__main__":
    print("hits", run(384, 1500, 300, 8, 2), run(384, 700, 40, 2, 2), run(768, 900, 140, 16, 4), run(64, 300, 7, 2, 2))
    # more work units than CTAs: the scans claim units dynamically (pair kernel, then streaming-A single-CTA)
    print("hits (claimed units)", run(384, 20000, 2100, 8, 2), run(768, 42000, 700, 8, 2))
    import torch
    w = synth.make_workload(2000, 8, 384, n_centroids=2, seed=3)
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    rows, ctx, ts, lru = t(w.rows), t(w.row_ctx.view(np.int64)), t(w.row_ts_us), t(w.row_lru.view(np.int64))
    q, qc = t(w.queries), t(w.q_ctx.view(np.int64))
    hr, hs = torch.zeros(8, dtype=torch.int64, device=dev), torch.zeros(8, dtype=torch.float32, device=dev)
    for _ in range(2):
        _cabi.exact_lookup_noalloc(0, 2000, 384, rows.data_ptr(), ctx.data_ptr(), ts.data_ptr(), lru.data_ptr(), 8,
                                   q.data_ptr(), qc.data_ptr(), 0.85, w.now_us, w.ttl_us, hr.data_ptr(), hs.data_ptr())
    torch.cuda.synchronize()
    print("noalloc", hr.cpu().tolist())
