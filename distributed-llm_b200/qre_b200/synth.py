"""Seeded synthetic workloads of the shapes BASELINE.json names.

Pure numpy, no dependency on the oracle or on CUDA: the same arrays feed the CUDA path,
the CPU oracle and the golden-fixture generator.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

HEURISTIC_PAIRS = (           # the fixed (device, confidence) set of qre:329-355
    (1, 0.92), (1, 0.80), (1, 0.88), (1, 0.75), (0, 0.90), (0, 0.75),
)


def _unit(x: np.ndarray) -> np.ndarray:
    return x / np.linalg.norm(x, axis=-1, keepdims=True)


def plant(e: np.ndarray, cos: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """Rows with a prescribed cosine to `e`: c*ê + sqrt(1-c²)*n̂, n̂ ⟂ ê (computed in fp64)."""
    e = _unit(e.astype(np.float64))
    n = rng.standard_normal(e.shape)
    n -= (n * e).sum(-1, keepdims=True) * e
    n = _unit(n)
    c = np.asarray(cos, np.float64)[..., None]
    return (c * e + np.sqrt(np.maximum(0.0, 1.0 - c * c)) * n)


@dataclass
class Workload:
    rows: np.ndarray            # float32 [N, d]
    row_ctx: np.ndarray         # uint64 [N]
    row_ts_us: np.ndarray       # int64 [N]
    row_lru: np.ndarray         # uint64 [N]
    row_pred_dev: np.ndarray    # uint8 [N]
    row_pred_conf: np.ndarray   # float64 [N]
    queries: np.ndarray         # float32 [M, d]
    q_ctx: np.ndarray           # uint64 [M]
    tok: np.ndarray             # int32 [M]
    heur_dev: np.ndarray        # uint8 [M]
    heur_conf: np.ndarray       # float64 [M]
    centroids: np.ndarray       # float32 [K, d]
    owner: np.ndarray           # uint8 [K]
    perf_health: np.ndarray     # float64 [D]
    now_us: int
    ttl_us: int
    planted_row: np.ndarray     # int64 [M]  (-1 = not planted) row the query was planted on
    planted_cos: np.ndarray     # float64 [M]


def ctx_id(i: int) -> int:
    """A 16-hex-digit context key as the u64 the device compares (router.py:145)."""
    return (0x9E3779B97F4A7C15 * (i + 1)) & 0xFFFFFFFFFFFFFFFE


def make_workload(n_rows: int, n_queries: int, dim: int = 384, n_centroids: int = 8,
                  n_devices: int = 2, seed: int = 0, threshold: float = 0.85,
                  frac_planted: float = 0.5, frac_near: float = 0.01, frac_dup: float = 0.001,
                  n_zero_q: int = 2, n_zero_rows: int = 2, frac_expired: float = 0.01,
                  n_ctx: int = 16, near_band: float = 1e-3,
                  rows: Optional[np.ndarray] = None) -> Workload:
    rs = np.random.Generator(np.random.PCG64(seed))
    if rows is None:
        rows = rs.standard_normal((n_rows, dim), dtype=np.float32)
        rows /= np.linalg.norm(rows, axis=1, keepdims=True)
        rows *= rs.uniform(0.5, 2.0, size=(n_rows, 1)).astype(np.float32)
    else:
        n_rows, dim = rows.shape
    # contexts: 90 % in context 0, the rest spread over n_ctx-1 others
    ctx_tab = np.array([ctx_id(i) for i in range(n_ctx)], np.uint64)
    row_ci = np.where(rs.random(n_rows) < 0.9, 0, rs.integers(1, n_ctx, n_rows))
    row_ctx = ctx_tab[row_ci]
    now_us = 1_800_000_000_000_000
    ttl_us = 3600 * 1_000_000
    age = rs.integers(0, ttl_us // 2, n_rows)
    expired = rs.random(n_rows) < frac_expired
    age = np.where(expired, ttl_us + 1 + rs.integers(0, 1000, n_rows), age)
    row_ts = (now_us - age).astype(np.int64)
    row_lru = rs.permutation(n_rows).astype(np.uint64) + np.uint64(1)
    row_pred_dev = rs.integers(0, n_devices, n_rows).astype(np.uint8)
    row_pred_conf = rs.uniform(0.5, 1.0, n_rows)
    if n_zero_rows and n_rows > n_zero_rows:
        zr = rs.choice(n_rows, n_zero_rows, replace=False)
        rows[zr] = 0.0

    rq = np.random.Generator(np.random.PCG64(seed + 1))
    Q = rq.standard_normal((n_queries, dim), dtype=np.float32)
    q_ci = np.where(rq.random(n_queries) < 0.9, 0, rq.integers(1, n_ctx, n_queries))
    planted_row = np.full(n_queries, -1, np.int64)
    planted_cos = np.zeros(n_queries)
    kind = rq.random(n_queries)
    live = np.nonzero(~expired & (np.abs(rows).sum(1) > 0))[0]
    for m in range(n_queries):
        if kind[m] < frac_planted + frac_near + frac_dup and live.size:
            j = int(live[rq.integers(0, live.size)])
            if kind[m] < frac_dup:
                c = 1.0
                Q[m] = rows[j] * np.float32(rq.uniform(0.5, 2.0))
            else:
                if kind[m] < frac_dup + frac_near:
                    c = threshold + rq.uniform(-near_band, near_band)
                else:
                    c = rq.uniform(0.86, 0.999)
                Q[m] = (plant(rows[j], np.array(c), rq) * rq.uniform(0.5, 2.0)).astype(np.float32)
            planted_row[m], planted_cos[m] = j, c
            q_ci[m] = row_ci[j]
    if n_zero_q:
        zq = rq.choice(n_queries, min(n_zero_q, n_queries), replace=False)
        Q[zq] = 0.0
        planted_row[zq] = -1
    q_ctx = ctx_tab[q_ci]

    rc = np.random.Generator(np.random.PCG64(seed + 2))
    dirs = _unit(rc.standard_normal((n_centroids, dim)))
    cents = np.stack([
        np.mean(_unit(dirs[k] + 0.6 * _unit(rc.standard_normal((30, dim)))).astype(np.float32),
                axis=0) for k in range(n_centroids)]).astype(np.float32)
    owner = (np.arange(n_centroids) % n_devices).astype(np.uint8)
    # pull ~35 % of the unplanted queries towards centroids so that every semantic branch
    # (clear win, ambiguous margin < 0.03, irrelevant < 0.05) is exercised
    for m in range(n_queries):
        if planted_row[m] >= 0 or not np.any(Q[m]):
            continue
        u = rc.random()
        if u < 0.25:
            k = rc.integers(0, n_centroids)
            Q[m] = (_unit(Q[m].astype(np.float64)) + rc.uniform(0.1, 0.8) * _unit(cents[k])
                    ).astype(np.float32)
        elif u < 0.35 and n_centroids >= 2:
            k0, k1 = rc.choice(n_centroids, 2, replace=False)
            if owner[k0] != owner[k1]:
                a = rc.uniform(0.2, 0.6)
                Q[m] = (_unit(Q[m].astype(np.float64)) + a * _unit(cents[k0])
                        + (a + rc.uniform(-0.02, 0.02)) * _unit(cents[k1])).astype(np.float32)

    rt = np.random.Generator(np.random.PCG64(seed + 3))
    tok = np.exp(rt.uniform(0.0, np.log(4000.0), n_queries)).astype(np.int32).clip(1)
    rh = np.random.Generator(np.random.PCG64(seed + 4))
    hp = rh.integers(0, len(HEURISTIC_PAIRS), n_queries)
    heur_dev = np.array([HEURISTIC_PAIRS[i][0] for i in hp], np.uint8) % n_devices
    heur_conf = np.array([HEURISTIC_PAIRS[i][1] for i in hp], np.float64)
    rp = np.random.Generator(np.random.PCG64(seed + 5))
    perf = rp.uniform(0.0, 1.0, n_devices)
    return Workload(rows, row_ctx, row_ts, row_lru, row_pred_dev, row_pred_conf,
                    Q, q_ctx, tok, heur_dev, heur_conf, cents, owner, perf,
                    now_us, ttl_us, planted_row, planted_cos)


def encoder_random_weights(hidden: int = 384, layers: int = 6, intermediate: int = 1536, vocab_size: int = 30522,
                           max_positions: int = 512, type_vocab: int = 2, seed: int = 0, std: float = 0.05) -> dict:
    """Seeded random parameters of a BERT-style encoder under transformers' `BertModel.state_dict()` names
    (the published all-MiniLM-L6-v2 checkpoint is not available offline): what the embedder benchmarks run on.
    LayerNorm gains near 1 and non-zero biases everywhere, so no term of a kernel is multiplied by zero."""
    rs = np.random.Generator(np.random.PCG64(seed))
    H, F = hidden, intermediate
    n = lambda *s: (rs.standard_normal(s) * std).astype(np.float32)   # noqa: E731
    g = lambda: (1.0 + 0.1 * rs.standard_normal(H)).astype(np.float32)   # noqa: E731
    w = {
        "embeddings.word_embeddings.weight": n(vocab_size, H),
        "embeddings.position_embeddings.weight": n(max_positions, H),
        "embeddings.token_type_embeddings.weight": n(type_vocab, H),
        "embeddings.LayerNorm.weight": g(),
        "embeddings.LayerNorm.bias": n(H),
    }
    for l in range(layers):
        p = f"encoder.layer.{l}."
        for nm in ("attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense"):
            w[p + nm + ".weight"] = n(H, H)
            w[p + nm + ".bias"] = n(H)
        w[p + "attention.output.LayerNorm.weight"] = g()
        w[p + "attention.output.LayerNorm.bias"] = n(H)
        w[p + "intermediate.dense.weight"] = n(F, H)
        w[p + "intermediate.dense.bias"] = n(F)
        w[p + "output.dense.weight"] = n(H, F)
        w[p + "output.dense.bias"] = n(H)
        w[p + "output.LayerNorm.weight"] = g()
        w[p + "output.LayerNorm.bias"] = n(H)
    return w
