#!/usr/bin/env python
"""Online-serving latency (BASELINE.json configs[4]): interleaved single-row cache inserts and
lookup+route batches of 1 / 8 / 64 queries on one B200, p50 / p99 of the host-visible latency
of the C-ABI calls (host buffers in, host results out), plus the HBM roofline of the scan at
small batch (algorithmic bytes = N * d * 2, the bf16 matrix read once)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "distributed-llm_b200"))
import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--entries", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=384)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--batches", default="1,8,64")
    a = ap.parse_args()
    import torch
    from qre_b200 import _cabi
    torch.cuda.set_device(0)
    args = argparse.Namespace(entries=a.entries, queries=4096, dim=a.dim, centroids=8, devices=2, keep_rows=False)
    wq = B.make_queries(args)
    sh, _ = B.build_shard(args, wq, 0, 1, torch)
    cfg = _cabi.RouteConfig(ttl_us=B.TTL_US)
    q_ctx = wq["ctx_tab"][wq["q_ci"]]
    rs = np.random.Generator(np.random.PCG64(7))
    pk = B.peaks()
    out = {"workload": f"interleaved 1 insert : 1 lookup+route batch, {a.entries} x {a.dim} cache, "
                       f"{a.steps} steps per batch size", "results": []}
    victim = np.arange(1000, 1000 + a.steps, dtype=np.int64)        # rows recycled by the inserts
    for mb in [int(x) for x in a.batches.split(",")]:
        lat_l, lat_i = [], []
        sh.profile(True)
        for i in range(a.steps + 50):
            j = (i * mb) % (4096 - mb)
            t0 = time.perf_counter()
            r = sh.lookup_route(wq["Q"][j:j + mb], q_ctx[j:j + mb], cfg, B.NOW_US, tok=wq["tok"][j:j + mb],
                                heur_dev=wq["heur_dev"][j:j + mb], heur_conf=wq["heur_conf"][j:j + mb])
            t1 = time.perf_counter()
            # insert = evict one row (tombstone) + append a fresh one into the freed slot
            v = victim[i % a.steps: i % a.steps + 1]
            if i == 0 or i >= a.steps:
                pass
            row = rs.standard_normal((1, a.dim)).astype(np.float32)
            t2 = time.perf_counter()
            sh.tombstone(v) if i < a.steps else None
            if i < a.steps:
                sh.append(row, q_ctx[:1], np.array([B.NOW_US], np.int64), np.array([10**9 + i], np.uint64),
                          np.array([1], np.uint8), np.array([0.9]))
            t3 = time.perf_counter()
            if i >= 50:
                lat_l.append(t1 - t0); lat_i.append(t3 - t2)
        ms, n = sh.profile_read()
        sh.profile(False)
        # the inserts landed in recycled slots 1000.. -> restore nothing; the cache stays 1M live rows
        scan_us = ms / max(n, 1) * 1e3
        bytes_alg = a.entries * a.dim * 2
        res = {"batch": mb,
               "lookup_route_p50_us": float(np.percentile(lat_l, 50) * 1e6),
               "lookup_route_p99_us": float(np.percentile(lat_l, 99) * 1e6),
               "insert_p50_us": float(np.percentile(lat_i[:a.steps - 50], 50) * 1e6),
               "insert_p99_us": float(np.percentile(lat_i[:a.steps - 50], 99) * 1e6),
               "routed_queries_per_s": mb / float(np.mean(lat_l)),
               "scan_kernel_us": scan_us,
               "roofline": {"bound": "hbm", "achieved": bytes_alg / (scan_us * 1e-6) / 1e9, "peak": pk["hbm"],
                            "unit": "GB/s", "frac": bytes_alg / (scan_us * 1e-6) / 1e9 / pk["hbm"],
                            "algorithmic_bytes": bytes_alg},
               "hits": int((r["hit_slot"] >= 0).sum())}
        out["results"].append(res)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
