"""CPU restatement of the reference's routing hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in the product (`distributed-llm_b200/`) may import this module: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs use it,
and only as the checker or the timed CPU baseline.

Parity pin: this restatement is checked against outputs of the reference's own code
(`/root/reference/src/cache.py`, `query_router_engine.py`) executed live in the build
container; the recorded outputs are committed under `tests/golden/` together with the
generator (`oracle/gen_golden.py`).  The reference itself ships no golden vectors (SURVEY
§4), so "the reference run here" is the pin.

All `file:line` citations are relative to /root/reference/src/.  `qre` = query_router_engine.py.

Two flavours of every scan are provided:
  * `*_faithful` — one Python iteration per cache entry using the same numpy scalar ops as
    the reference (`np.dot`, `np.linalg.norm`), i.e. bit-identical to the reference when the
    numpy/BLAS build is the same.  This is the CPU baseline `bench.py` times.
  * `*_batch`    — vectorised (sgemm) form used as the fast oracle at 1M rows.  Scores differ
    from the faithful form only by fp32 summation order (≈1e-7 relative).

Generalisation to K centroids / D devices follows SURVEY §8(a) notes and reduces to the
reference at K = D = 2, w_perf = 0 (device 0 = "nano", device 1 = "orin").
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

RECENCY_DECAY = 0.85          # cache.py:46
MAX_HISTORY = 20              # cache.py:92
NORM_EPS = 1e-9               # cache.py:284,289 ; qre:136
CTX_TOMBSTONE = np.uint64(0xFFFFFFFFFFFFFFFF)

# method codes shared with the CUDA epilogue (include/qre_b200.h)
SEM_METHOD_SEMANTIC = 0       # qre:207-213
SEM_METHOD_IRRELEVANT = 1     # qre:185-193
SEM_METHOD_AMBIGUOUS = 2      # qre:195-204
SEM_METHOD_NAMES = ("semantic", "semantic_fallback_irrelevant", "semantic_fallback_ambiguous")


# ----------------------------------------------------------------------------------------
# Scalar arithmetic, restated with the reference's rounding sequence
# ----------------------------------------------------------------------------------------

def cache_similarity(q: np.ndarray, e: np.ndarray) -> Optional[float]:
    """cache.py:283-291.  norms are fp32 `sqrt(sdot)` widened to Python float; the product
    of the two is exact in fp64; `np.float32 / python float` is an fp32 divide (numpy 2
    weak-scalar promotion), so sim = fl32(sdot32(q,e) / fl32(nq*ne)).  Returns None where
    the reference `continue`s / returns (norm < 1e-9)."""
    nq = float(np.linalg.norm(q))
    if nq < NORM_EPS:
        return None
    ne = float(np.linalg.norm(e))
    if ne < NORM_EPS:
        return None
    return float(np.dot(q, e) / (nq * ne))


def centroid_similarity(a: np.ndarray, b: np.ndarray) -> float:
    """qre:133-138.  All fp32: sdot32 / (||a||32 * ||b||32); 0.0 when a norm < 1e-9."""
    na, nb = np.linalg.norm(a), np.linalg.norm(b)
    if na < NORM_EPS or nb < NORM_EPS:
        return 0.0
    return float(np.dot(a, b) / (na * nb))


def token_decision(tok: int, threshold: int) -> Tuple[int, float]:
    """qre:98-99.  device 1 ("orin") iff tok > threshold; conf = min(|tok-thr|/max(thr,1), 1)."""
    dev = 1 if tok > threshold else 0
    conf = float(min(abs(tok - threshold) / max(threshold, 1), 1.0))
    return dev, conf


def approx_token_count(query: str, context: Optional[str]) -> int:
    """qre:91,96 — the reference's own fallback when litellm is absent."""
    full = f"{context}\n{query}" if context else query
    return max(1, len(full) // 4)


def predict_device(history: Sequence[Tuple[int, float]], device_used: int,
                   n_devices: int = 2) -> Tuple[int, float]:
    """cache.py:106-140.  `history` is oldest-first [(device, confidence)].  Newest-first
    weights 0.85**i * conf; winner share = score/total; ties -> HIGHEST index (cache.py:133
    `orin_score >= nano_score`); empty history or ~zero weight -> (device_used, 0.5)."""
    if not history:
        return device_used, 0.5
    scores = [0.0] * n_devices
    total = 0.0
    for i, (dev, conf) in enumerate(reversed(list(history))):
        w = (RECENCY_DECAY ** i) * conf
        scores[dev] += w
        total += w
    if total < 1e-9:
        return device_used, 0.5
    best = 0
    for d in range(1, n_devices):
        if scores[d] >= scores[best]:
            best = d
    return best, float(min(scores[best] / total, 1.0))


@dataclass
class SemanticParams:
    min_similarity: float = 0.05      # qre:129
    margin_threshold: float = 0.03    # qre:128


def semantic_decision(sims: Sequence[float], owner: Sequence[int], n_devices: int,
                      tok_dev: int, tok_conf: float,
                      p: SemanticParams = SemanticParams()) -> Tuple[int, float, int, List[float]]:
    """qre:180-213, generalised.  `sims[k]` are the fp32 centroid cosines widened to float,
    `owner[k]` the device that owns centroid k.  Sim[dev] = max over owned centroids.
    Returns (device, confidence, method_code, Sim[dev] list).

    K = D = 2, owner = (0, 1) reproduces: irrelevant iff both sims < min_similarity
    (qre:185); margin = |sim_orin - sim_nano| (qre:195); ambiguous iff margin < 0.03
    (qre:196); else orin iff sim_orin > sim_nano, tie -> nano (qre:206);
    conf = min(1, margin/0.2) (qre:209)."""
    dev_sim = [-math.inf] * n_devices
    for s, o in zip(sims, owner):
        if s > dev_sim[o]:
            dev_sim[o] = s
    have = [d for d in range(n_devices) if dev_sim[d] != -math.inf]
    if all(dev_sim[d] < p.min_similarity for d in have):
        return tok_dev, tok_conf * 0.5, SEM_METHOD_IRRELEVANT, dev_sim
    top = have[0]
    for d in have[1:]:
        if dev_sim[d] > dev_sim[top]:
            top = d
    rest = [dev_sim[d] for d in have if d != top]
    second = max(rest) if rest else -math.inf
    margin = abs(dev_sim[top] - second) if rest else math.inf
    if margin < p.margin_threshold:
        return tok_dev, float(margin), SEM_METHOD_AMBIGUOUS, dev_sim
    return top, float(min(1.0, margin / 0.2)), SEM_METHOD_SEMANTIC, dev_sim


def hybrid_decision(votes: Sequence[Tuple[int, float, float]], n_devices: int = 2,
                    w_perf: float = 0.0, perf_health: Optional[Sequence[float]] = None
                    ) -> Tuple[int, float, List[float]]:
    """qre:382-414, generalised.  `votes` = [(device, confidence, weight)] in the order
    token, semantic, heuristic (dict order, qre:376-380).  S[dev] += w*conf.  Optional
    added term w_perf * H[dev] (absent from the reference; skipped when w_perf == 0).
    Winner = argmax S, ties -> LOWEST index (qre:401 strict `>`); margin = S[top]-S[2nd];
    total = sum S in index order; conf = clamp(margin/total if total > 1e-12 else 0.5)."""
    S = [0.0] * n_devices
    for dev, conf, w in votes:
        S[dev] += float(w) * float(conf)
    if w_perf != 0.0 and perf_health is not None:
        for d in range(n_devices):
            S[d] += w_perf * float(perf_health[d])
    top = 0
    for d in range(1, n_devices):
        if S[d] > S[top]:
            top = d
    second = max(S[d] for d in range(n_devices) if d != top)
    margin = S[top] - second
    total = 0.0
    for d in range(n_devices):
        total += S[d]
    conf = float(min(max(margin / total if total > 1e-12 else 0.5, 0.0), 1.0))
    return top, conf, S


def perf_score(samples: Sequence[Tuple[float, int, int]], fail_penalty: float = 3000.0) -> float:
    """qre:435-445.  samples = [(latency_ms, tokens, ok)] within the window."""
    if not samples:
        return math.inf
    total_lat = sum(x[0] for x in samples)
    total_tok = sum(x[1] for x in samples)
    ok_sum = sum(x[2] for x in samples)
    fail_rate = 1.0 - (ok_sum / len(samples))
    if total_tok == 0:
        return float(total_lat / len(samples) + fail_penalty * fail_rate)
    return float(total_lat / total_tok + fail_penalty * fail_rate)


def perf_decision(scores: Sequence[float]) -> Tuple[int, float]:
    """qre:447-458, generalised: lower score wins, ties -> lowest index; all-inf -> (0, 0.2)."""
    if all(s == math.inf for s in scores):
        return 0, 0.2
    best = 0
    for d in range(1, len(scores)):
        if scores[d] < scores[best]:
            best = d
    return best, 0.70


# ----------------------------------------------------------------------------------------
# Struct-of-arrays cache image (what the device holds) and the scans over it
# ----------------------------------------------------------------------------------------

@dataclass
class CacheImage:
    """Rows of the routing cache as flat arrays.  Mirrors the per-row state the reference
    keeps in `CacheEntry` (cache.py:79-89) that the semantic scan reads:
    embedding, context_key, timestamp, plus the OrderedDict position (`lru_seq`)."""
    rows: np.ndarray          # float32 [N, d]
    ctx: np.ndarray           # uint64  [N]   (CTX_TOMBSTONE = deleted / no embedding)
    ts_us: np.ndarray         # int64   [N]   microseconds since epoch
    lru_seq: np.ndarray       # uint64  [N]   smaller = earlier in OrderedDict order
    pred_dev: np.ndarray = field(default=None)    # uint8 [N]  predict_device() result
    pred_conf: np.ndarray = field(default=None)   # float64 [N]

    @property
    def n(self) -> int:
        return int(self.rows.shape[0])


def _row_mask(img: CacheImage, ctx: int, now_us: int, ttl_us: int) -> np.ndarray:
    """cache.py:274-276: same context, embedding present, not expired
    ((now - ts).total_seconds() <= ttl  <=>  now_us - ts_us <= ttl_us, exact in µs)."""
    return (img.ctx == np.uint64(ctx)) & (img.ctx != CTX_TOMBSTONE) & \
           ((np.int64(now_us) - img.ts_us) <= np.int64(ttl_us))


def scan_faithful(q: np.ndarray, ctx: int, img: CacheImage, now_us: int, ttl_us: int,
                  threshold: float) -> Tuple[int, float]:
    """cache.py:280-296.  Returns (row, sim) of the hit or (-1, 0.0) on a miss.  Candidates
    are visited in OrderedDict order (ascending lru_seq); only a strictly greater sim
    replaces the best (cache.py:292), and best_sim starts at 0.0 so sims <= 0 never hit."""
    best_row, best_sim = -1, 0.0
    norm_q = float(np.linalg.norm(q))
    if norm_q < NORM_EPS:
        return -1, 0.0
    mask = _row_mask(img, ctx, now_us, ttl_us)
    order = np.argsort(img.lru_seq, kind="stable")
    for j in order:
        if not mask[j]:
            continue
        e = img.rows[j]
        norm_c = float(np.linalg.norm(e))
        if norm_c < NORM_EPS:
            continue
        sim = float(np.dot(q, e) / (norm_q * norm_c))
        if sim >= threshold and sim > best_sim:
            best_sim, best_row = sim, int(j)
    return best_row, best_sim


def row_norms32(x: np.ndarray) -> np.ndarray:
    """fp32 ||row|| = sqrt32(sum32(x*x)) (np.linalg.norm on fp32 stays fp32)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    return np.sqrt(np.einsum("ij,ij->i", x, x, dtype=np.float32), dtype=np.float32)


def sims_batch(Q: np.ndarray, E: np.ndarray, nq: Optional[np.ndarray] = None,
               ne: Optional[np.ndarray] = None) -> np.ndarray:
    """Vectorised cache.py:291: fl32(sdot32 / fl32(nq*ne)) for every (query, row) pair.
    Rows/queries with norm < 1e-9 get -inf (never a hit)."""
    nq = row_norms32(Q) if nq is None else nq
    ne = row_norms32(E) if ne is None else ne
    dots = np.asarray(Q, np.float32) @ np.asarray(E, np.float32).T
    den = (nq.astype(np.float64)[:, None] * ne.astype(np.float64)[None, :]).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        s = dots / den
    s[nq < NORM_EPS, :] = -np.inf
    s[:, ne < NORM_EPS] = -np.inf
    return s


def scan_batch(Q: np.ndarray, q_ctx: np.ndarray, img: CacheImage, now_us: int, ttl_us: int,
               threshold: float, chunk: int = 65536, topk: int = 1):
    """Batched form of cache.py:280-296 against one snapshot of the cache.

    Returns (hit_row int64[M], hit_sim float32[M]) when topk == 1; with topk > 1 also returns
    (top_rows int64[M,topk], top_sims float32[M,topk]) ordered by (sim desc, lru_seq asc),
    restricted to sims that pass `sim >= threshold and sim > 0`."""
    Q = np.ascontiguousarray(Q, np.float32)
    M = Q.shape[0]
    nq = row_norms32(Q)
    best_sim = np.full(M, -np.inf, np.float32)
    best_lru = np.full(M, np.iinfo(np.uint64).max, np.uint64)
    best_row = np.full(M, -1, np.int64)
    keep_rows: List[np.ndarray] = []
    keep_sims: List[np.ndarray] = []
    valid_row = (img.ctx != CTX_TOMBSTONE) & ((np.int64(now_us) - img.ts_us) <= np.int64(ttl_us))
    thr32_ok = np.float64(threshold)
    for s0 in range(0, img.n, chunk):
        s1 = min(img.n, s0 + chunk)
        S = sims_batch(Q, img.rows[s0:s1], nq=nq)
        ok = (q_ctx[:, None] == img.ctx[None, s0:s1]) & valid_row[None, s0:s1]
        ok &= (S.astype(np.float64) >= thr32_ok) & (S > 0)
        S = np.where(ok, S, -np.inf).astype(np.float32)
        if topk > 1:
            keep_rows.append(np.arange(s0, s1))
            keep_sims.append(S)
        cmax = S.max(axis=1)
        lru = img.lru_seq[s0:s1]
        # among rows equal to the chunk max pick the smallest lru_seq
        is_max = (S == cmax[:, None]) & np.isfinite(cmax)[:, None]
        lru_m = np.where(is_max, lru[None, :], np.iinfo(np.uint64).max)
        arg = lru_m.argmin(axis=1)
        c_lru = lru_m[np.arange(M), arg]
        better = (cmax > best_sim) | ((cmax == best_sim) & (c_lru < best_lru))
        better &= np.isfinite(cmax)
        best_sim = np.where(better, cmax, best_sim)
        best_lru = np.where(better, c_lru, best_lru)
        best_row = np.where(better, arg + s0, best_row)
    hit_sim = np.where(best_row >= 0, best_sim, np.float32(0.0)).astype(np.float32)
    if topk == 1:
        return best_row, hit_sim
    allS = np.concatenate(keep_sims, axis=1)
    lru_all = img.lru_seq.astype(np.float64)  # only used as a sort key
    top_rows = np.full((M, topk), -1, np.int64)
    top_sims = np.full((M, topk), -np.inf, np.float32)
    for m in range(M):
        idx = np.nonzero(np.isfinite(allS[m]))[0]
        if idx.size == 0:
            continue
        order = np.lexsort((img.lru_seq[idx], -allS[m, idx].astype(np.float64)))[:topk]
        sel = idx[order]
        top_rows[m, :sel.size] = sel
        top_sims[m, :sel.size] = allS[m, sel]
    del lru_all
    return best_row, hit_sim, top_rows, top_sims


def centroid_sims_batch(Q: np.ndarray, C: np.ndarray) -> np.ndarray:
    """qre:133-138 for every (query, centroid): fp32 dot / (fp32 nq * fp32 nc); 0 on tiny norm."""
    Q = np.ascontiguousarray(Q, np.float32)
    C = np.ascontiguousarray(C, np.float32)
    nq, nc = row_norms32(Q), row_norms32(C)
    den = nq[:, None] * nc[None, :]
    with np.errstate(divide="ignore", invalid="ignore"):
        s = (Q @ C.T) / den
    s[(nq < NORM_EPS)[:, None] | (nc < NORM_EPS)[None, :]] = 0.0
    return s.astype(np.float32)


@dataclass
class RouteParams:
    """Knobs of the fused route (defaults = shipped configs, qre:704-731)."""
    n_devices: int = 2
    token_threshold: int = 1000                 # qre:522
    w_token: float = 0.25                       # qre:539
    w_semantic: float = 0.45
    w_heuristic: float = 0.30
    w_perf: float = 0.0                         # added term, 0 in the reference
    sem: SemanticParams = field(default_factory=SemanticParams)
    similarity_threshold: float = 0.85          # qre:500
    prediction_confidence_threshold: float = 0.70  # qre:502-504
    ttl_us: int = 3600 * 1_000_000              # qre:499


def route_batch(Q, tok, heur_dev, heur_conf, C, owner, p: RouteParams, perf_health=None):
    """Batched HybridRouter.route (qre:382-414) over precomputed host features.
    Returns dict of per-query arrays (sims, sem_dev/conf/method, tok_dev/conf, dev, conf, S)."""
    M = Q.shape[0]
    sims = centroid_sims_batch(Q, C)
    out = dict(
        sims=sims,
        tok_dev=np.zeros(M, np.uint8), tok_conf=np.zeros(M, np.float64),
        sem_dev=np.zeros(M, np.uint8), sem_conf=np.zeros(M, np.float64),
        sem_method=np.zeros(M, np.uint8),
        dev=np.zeros(M, np.uint8), conf=np.zeros(M, np.float64),
        S=np.zeros((M, p.n_devices), np.float64),
    )
    for m in range(M):
        td, tc = token_decision(int(tok[m]), p.token_threshold)
        sd, sc, sm, _ = semantic_decision([float(x) for x in sims[m]], owner, p.n_devices,
                                          td, tc, p.sem)
        votes = [(td, tc, p.w_token), (sd, sc, p.w_semantic),
                 (int(heur_dev[m]), float(heur_conf[m]), p.w_heuristic)]
        d, c, S = hybrid_decision(votes, p.n_devices, p.w_perf, perf_health)
        out["tok_dev"][m], out["tok_conf"][m] = td, tc
        out["sem_dev"][m], out["sem_conf"][m], out["sem_method"][m] = sd, sc, sm
        out["dev"][m], out["conf"][m], out["S"][m] = d, c, S
    return out
