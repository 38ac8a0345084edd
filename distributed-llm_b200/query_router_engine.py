"""query_router_engine.py — drop-in for the reference's `src/query_router_engine.py`.

Same classes, constructor arguments, config keys and `RoutingDecision` fields.  What moves to
the B200 (libqre_b200.so, see include/qre_b200.h):

  * SemanticRouter's centroid cosines and its three-way rule (reference qre:133-138, 180-213);
  * TokenBasedRouter's compare + confidence (qre:98-99) and HybridRouter's weighted vote and
    argmax (qre:382-414) as the epilogue of the same kernel sequence;
  * together with the cache's semantic scan (`cache.QueryCache`), so one `QueryRouter.
    route_query` call is ONE device pass: lookup + centroid scores + hybrid vote.

What stays on the host is string work: token counting, the heuristic regexes (qre:231-364),
the reasoning strings, and the hit / override / insert control flow (qre:575-645).

Generalisation beyond the reference: K centroids with an owner device each, D
devices (`config["devices"]`), and an optional perf term `w_perf * H[dev]` in the hybrid
score; with the default config (two centroids, nano/orin, w_perf = 0) every output is the
reference's.
"""
from __future__ import annotations

import json
import logging
import os
import re
from abc import ABC, abstractmethod
from collections import deque
from dataclasses import dataclass
from datetime import datetime
from typing import Any, Dict, List, Literal, Optional, Sequence, Tuple

import numpy as np

try:
    from cache import CacheEntry, CacheLookupResult, QueryCache, DEVICE_NAMES, _gc_paused
except ImportError as _e:   # pragma: no cover
    raise ImportError("query_router_engine needs the B200 cache.py next to it on sys.path "
                      f"(distributed-llm_b200/): {_e}") from _e

from qre_b200 import _cabi, heuristic_tables
from qre_b200.engine import Engine, ShardedEngine


def _new_engine(config, capacity: int, max_batch: int = 64):
    """Device store from a config dict: `gpus` = [i, j, ...] shards the rows over those GPUs (one
    process drives them all), `gpu` = i (default 0) keeps everything on one."""
    cfg = config or {}
    gpus = list(cfg["gpus"]) if cfg.get("gpus") is not None else []
    if len(gpus) > 1:
        return ShardedEngine(capacity=capacity, devices=gpus, max_batch=max_batch)
    return Engine(capacity=capacity, device=int(gpus[0]) if gpus else int(cfg.get("gpu", 0)), max_batch=max_batch)


logger = logging.getLogger(__name__)

try:
    from sentence_transformers import SentenceTransformer
    SENTENCE_TRANSFORMERS_AVAILABLE = True
except ImportError:
    SENTENCE_TRANSFORMERS_AVAILABLE = False
    logger.warning("sentence-transformers not available. Semantic routing disabled.")

try:
    from litellm import token_counter
    LITELLM_AVAILABLE = True
except ImportError:
    LITELLM_AVAILABLE = False
    logger.warning("litellm not available. Token-based routing will use approximation.")


@dataclass
class RoutingDecision:
    device: Literal["nano", "orin"]
    confidence: float
    method: str
    reasoning: str
    complexity_score: Optional[float] = None
    cache_hit: bool = False


class LazyRoutingDecision(RoutingDecision):
    """A RoutingDecision whose `reasoning` string is formatted on first access (batch path: 4096
    f-strings per batch cost more than the device pass).  Everything else — fields, equality with
    plain decisions, repr — is the reference's record (qre:55-62)."""

    def __init__(self, device, confidence, method, reasoning_fn, reasoning_arg=None, complexity_score=None,
                 cache_hit=False):
        self.__dict__.update(device=device, confidence=confidence, method=method,
                             complexity_score=complexity_score, cache_hit=cache_hit,
                             _lazy=(reasoning_fn, reasoning_arg))

    def __getattr__(self, name):
        if name == "reasoning":
            fn, arg = self.__dict__.pop("_lazy")
            self.__dict__["reasoning"] = r = fn(arg)
            return r
        raise AttributeError(name)

    def _t(self):
        return (self.device, self.confidence, self.method, self.reasoning, self.complexity_score, self.cache_hit)

    def __eq__(self, other):
        return isinstance(other, RoutingDecision) and self._t() == (
            other.device, other.confidence, other.method, other.reasoning, other.complexity_score, other.cache_hit)

    __hash__ = None

    def __repr__(self):
        return ("RoutingDecision(device={!r}, confidence={!r}, method={!r}, reasoning={!r}, "
                "complexity_score={!r}, cache_hit={!r})").format(*self._t())


def _cached_why(t) -> str:
    """Reasoning string of a cached decision from its snapshot tuple (reference qre:585-592)."""
    return (f"cache hit age={int((t[0] - t[1]).total_seconds())}s hits={t[2]} predicted={t[3]} conf={t[4]:.2f} "
            f"context_len={t[5]} history={t[6]}")


class BaseRouter(ABC):
    def __init__(self, config: Dict[str, Any]):
        self.config = config

    @abstractmethod
    def route(self, query: str, context: Optional[str] = None) -> RoutingDecision:
        raise NotImplementedError


def _device_names(config: Dict[str, Any]) -> List[str]:
    return list(config.get("devices", DEVICE_NAMES))


# ----------------------------------------------------------------------------- token
class TokenBasedRouter(BaseRouter):
    """device = the heavy one iff tokens > threshold (reference qre:82-107)."""

    def __init__(self, config: Dict[str, Any]):
        super().__init__(config)
        self.threshold = int(config.get("token_threshold", 1000))
        self.model = config.get("model", "meta-llama/Llama-2-7b-hf")
        self._names = _device_names(config)

    def count(self, query: str, context: Optional[str]) -> int:
        text = f"{context}\n{query}" if context else query
        if LITELLM_AVAILABLE:
            return int(token_counter(model=self.model, text=text))
        return max(1, len(text) // 4)

    def decide(self, tok: int) -> Tuple[int, float]:
        """The arithmetic the GPU epilogue repeats (qre:98-99)."""
        return (1 if tok > self.threshold else 0,
                float(min(abs(tok - self.threshold) / max(self.threshold, 1), 1.0)))

    def _decision(self, tok: int, dev: int, conf: float) -> RoutingDecision:
        return RoutingDecision(device=self._names[dev], confidence=conf, method="token",
                               reasoning=f"tokens={tok} threshold={self.threshold}",
                               complexity_score=float(tok))

    def route(self, query: str, context: Optional[str] = None) -> RoutingDecision:
        tok = self.count(query, context)
        return self._decision(tok, *self.decide(tok))


def _make_embedder(config: Dict[str, Any]):
    """The object whose `.encode(list[str]) -> float32 [n, d]` the routers call (qre:126, 511).
    `config["embedder"]` (additive key) supplies one directly — e.g. a
    `qre_b200.embedder.B200SentenceEncoder`, which embeds on the GPU next to the cache; otherwise the
    reference's SentenceTransformer(model name) when that package is installed, else None."""
    emb = config.get("embedder")
    if emb is not None:
        return emb
    if SENTENCE_TRANSFORMERS_AVAILABLE:
        return SentenceTransformer(config.get("embedding_model", "all-MiniLM-L6-v2"))
    return None


# ----------------------------------------------------------------------------- semantic
class SemanticRouter(BaseRouter):
    """Centroid router (reference qre:114-213); the cosines and the rule run on the GPU."""

    def __init__(self, config: Dict[str, Any], engine: Optional[Engine] = None):
        super().__init__(config)
        self.embedder = _make_embedder(config)
        if self.embedder is None:
            raise ImportError("sentence-transformers required for SemanticRouter")
        self.label_path = config.get("semantic_label_path", "")
        self.margin_threshold = float(config.get("semantic_margin_threshold", 0.03))
        self.min_similarity = float(config.get("semantic_min_similarity", 0.05))
        self._token_fallback = TokenBasedRouter(config)
        self._names = _device_names(config)
        if "centroids" in config:                      # generalised: explicit K x d matrix + owners
            self.centroids = np.ascontiguousarray(config["centroids"], np.float32)
            self.owner = np.ascontiguousarray(config["centroid_owner"], np.uint8)
        else:
            self.nano_center, self.orin_center = self._load_and_build_centroids(self.label_path)
            self.centroids = np.stack([self.nano_center, self.orin_center]).astype(np.float32)
            self.owner = np.array([0, 1], np.uint8)
        self._engine = engine if engine is not None else _new_engine(config, 1)
        self._engine.set_centroids(self.centroids, self.owner)

    _SIMPLE_SEED = ("Hello", "What is 2+2?", "Define machine learning", "What is the weather today?")
    _COMPLEX_SEED = (
        "Write a Python function to solve knapsack and explain time complexity",
        "Analyze the economic impact of climate policy with trade-offs",
        "Draft a detailed report with methodology and evaluation plan",
        "Explain quantum computing implications for cryptography in depth",
    )

    def _load_and_build_centroids(self, label_path: str) -> Tuple[np.ndarray, np.ndarray]:
        """fp32 mean of the label embeddings per class (reference qre:140-178)."""
        if not label_path or not os.path.exists(label_path):
            groups = {"nano": list(self._SIMPLE_SEED), "orin": list(self._COMPLEX_SEED)}
        else:
            with open(label_path, "r", encoding="utf-8") as f:
                rows = json.load(f)
            groups = {"nano": [], "orin": []}
            for row in rows:
                text = (row.get("text") or "").strip()
                label = (row.get("label") or "").strip().lower()
                if text and label in groups:
                    groups[label].append(text)
            if len(groups["nano"]) < 3 or len(groups["orin"]) < 3:
                raise ValueError(f"Need >=3 samples per class. Got nano={len(groups['nano'])} "
                                 f"orin={len(groups['orin'])}")
        return (np.mean(self.embedder.encode(groups["nano"]), axis=0),
                np.mean(self.embedder.encode(groups["orin"]), axis=0))

    def _route_cfg(self) -> _cabi.RouteConfig:
        c = self.config
        w = c.get("weights", {"token": 0.35, "semantic": 0.35, "heuristic": 0.30})
        return _cabi.RouteConfig(
            n_devices=len(self._names), token_threshold=self._token_fallback.threshold,
            w_token=float(w.get("token", 0.0)), w_semantic=float(w.get("semantic", 0.0)),
            w_heuristic=float(w.get("heuristic", 0.0)), w_perf=float(c.get("perf_weight", 0.0)),
            sem_min_similarity=self.min_similarity, sem_margin_threshold=self.margin_threshold)

    def _device_pass(self, q_emb: np.ndarray, tok: int, heur: Tuple[int, float] = (0, 0.0)) -> dict:
        route_in = dict(tok=np.array([tok], np.int32), heur_dev=np.array([heur[0]], np.uint8),
                        heur_conf=np.array([heur[1]], np.float64))
        return self._engine.scan(q_emb, np.array([0], np.uint64), 0, self._route_cfg(), route_in)

    def _decision(self, out: dict, i: int, tok_decision: RoutingDecision) -> RoutingDecision:
        """Rebuild the reference's RoutingDecision (qre:185-213) from the kernel outputs."""
        sims = out["centroid_sim"][i]
        dsim = [max([float(s) for s, o in zip(sims, self.owner) if o == d], default=float("-inf"))
                for d in range(len(self._names))]
        sim_n, sim_o = dsim[0], dsim[-1]
        method = _cabi.SEM_METHOD_NAMES[int(out["sem_method"][i])]
        conf = float(out["sem_conf"][i])
        dev = self._names[int(out["sem_dev"][i])]
        if method == "semantic_fallback_irrelevant":
            why = f"low similarity (n={sim_n:.2f}, o={sim_o:.2f}) -> {tok_decision.reasoning}"
        elif method == "semantic_fallback_ambiguous":
            why = f"ambiguous margin={conf:.3f} (n={sim_n:.2f}, o={sim_o:.2f}) -> {tok_decision.reasoning}"
        else:
            top2 = sorted(dsim, reverse=True)[:2]
            margin = abs(top2[0] - top2[1]) if len(top2) > 1 else 0.0
            why = f"sim_nano={sim_n:.3f} sim_orin={sim_o:.3f} margin={margin:.3f}"
        return RoutingDecision(device=dev, confidence=conf, method=method, reasoning=why,
                               complexity_score=float(sim_o))

    def route(self, query: str, context: Optional[str] = None) -> RoutingDecision:
        q_emb = self.embedder.encode([query])[0]
        tok_d = self._token_fallback.route(query, context)
        out = self._device_pass(q_emb, int(tok_d.complexity_score))
        return self._decision(out, 0, tok_d)


# ----------------------------------------------------------------------------- heuristic
class HeuristicRouter(BaseRouter):
    """Rule-based router (reference qre:220-364).  Host-side string work; its (device,
    confidence) pair enters the GPU epilogue as a per-query input.  The pattern tables are the
    reference's (they ARE the behaviour); the fixed confidences are 0.92 / 0.80 / 0.88 / 0.75
    for the heavy device and 0.90 / 0.75 for the light one."""

    COMPLEX = {name: list(pats) for name, pats in heuristic_tables.COMPLEX_CATEGORIES}
    SIMPLE = {name: list(pats) for name, pats in heuristic_tables.SIMPLE_CATEGORIES}
    CODE_MARKERS = heuristic_tables.CODE_MARKERS

    def __init__(self, config: Dict[str, Any]):
        super().__init__(config)
        self.long_text_threshold = int(config.get("heuristic_long_chars", 250))
        self.multi_question_threshold = int(config.get("heuristic_multi_qmarks", 3))
        self.code_markers_needed = int(config.get("heuristic_code_markers_needed", 2))
        self.context_chars_threshold = int(config.get("heuristic_context_chars", 800))
        self._token_fallback = TokenBasedRouter(config)
        self._names = _device_names(config)
        comp = lambda table: {k: [re.compile(p, re.IGNORECASE) for p in v] for k, v in table.items()}  # noqa: E731
        self.complex_patterns = comp(self.COMPLEX)
        self.simple_patterns = comp(self.SIMPLE)

    @staticmethod
    def _first_category(text: str, table: Dict[str, list]) -> Optional[str]:
        for cat, pats in table.items():
            if any(p.search(text) for p in pats):
                return cat
        return None

    def classify(self, query: str, context: Optional[str]) -> Tuple[int, float, str, str, Optional[float]]:
        """-> (device index, confidence, method, reasoning, complexity_score); the order of the
        rules is the reference's (qre:323-364)."""
        q = (query or "").strip()
        low = q.lower()
        heavy, light = len(self._names) - 1, 0
        cat = self._first_category(low, self.complex_patterns)
        if cat is not None:
            return heavy, 0.92, "heuristic", f"complex pattern={cat}", None
        if len(q) >= self.long_text_threshold:
            return heavy, 0.80, "heuristic", f"long query chars={len(q)}", None
        if q.count("?") >= self.multi_question_threshold:
            return heavy, 0.80, "heuristic", f"multi-question count={q.count('?')}", None
        if sum(1 for mk in self.CODE_MARKERS if mk in q) >= self.code_markers_needed:
            return heavy, 0.88, "heuristic", "code/debug markers detected", None
        if context and len(context) >= self.context_chars_threshold:
            return heavy, 0.75, "heuristic", f"large context chars={len(context)}", None
        scat = self._first_category(low, self.simple_patterns)
        if scat is not None:
            return light, 0.90, "heuristic", f"simple pattern={scat}", None
        if len(low.split()) <= 15 and len(q) <= 100:
            return light, 0.75, "heuristic", "short everyday query", None
        d = self._token_fallback.route(query, context)
        return (self._names.index(d.device), float(d.confidence * 0.5), "heuristic_fallback",
                f"no heuristic match -> {d.reasoning}", d.complexity_score)

    def route(self, query: str, context: Optional[str] = None) -> RoutingDecision:
        dev, conf, method, why, score = self.classify(query, context)
        return RoutingDecision(device=self._names[dev], confidence=conf, method=method, reasoning=why,
                               complexity_score=score)

    def features_batch(self, queries: Sequence[str], contexts: Sequence[Optional[str]],
                       context_keys: Optional[Sequence[str]] = None):
        """What the fused device pass needs from the host for a batch — token counts and the
        heuristic (device, confidence) pairs — as arrays, without building any per-query object
        (qre_host_features_packed on C++ threads; the rare query it hands back, and every query when a real
        tokenizer is installed, go through the Python rules).  With `context_keys` a fourth value is
        returned: the cache's exact-match key of every query (computed in the same C++ pass)."""
        if LITELLM_AVAILABLE:
            feats, toks = self.classify_batch(queries, contexts)
            r3 = (np.asarray(toks, np.int32), np.array([f[0] for f in feats], np.uint8),
                  np.array([f[1] for f in feats], np.float64))
            return r3 if context_keys is None else r3 + ([None] * len(queries),)
        hp = _cabi.HeurParams(self.long_text_threshold, self.multi_question_threshold, self.code_markers_needed,
                              self.context_chars_threshold, self._token_fallback.threshold, len(self._names))
        r = _cabi.host_features_packed([(q or "") for q in queries], [len(c) if c else 0 for c in contexts], hp,
                                       context_keys=context_keys)
        tok, hd, hc = r["tok"], r["heur_dev"], r["heur_conf"]
        for i in np.nonzero(r["rule"] < 0)[0].tolist():
            f = self.classify(queries[i], contexts[i])
            hd[i], hc[i] = f[0], f[1]
            tok[i] = self._token_fallback.count(queries[i], contexts[i])
        return (tok, hd, hc) if context_keys is None else (tok, hd, hc, r["keys"])

    def classify_batch(self, queries: Sequence[str], contexts: Sequence[Optional[str]]):
        """`classify` for a whole batch on C++ threads (libqre_b200 qre_host_features); the few
        queries it hands back (none for well-formed text), and every query when a real tokenizer is installed,
        go through `classify`.  Returns (list of classify() tuples, token counts)."""
        n = len(queries)
        if LITELLM_AVAILABLE:
            feats = [self.classify(q, c) for q, c in zip(queries, contexts)]
            return feats, [self._token_fallback.count(q, c) for q, c in zip(queries, contexts)]
        hp = _cabi.HeurParams(self.long_text_threshold, self.multi_question_threshold, self.code_markers_needed,
                              self.context_chars_threshold, self._token_fallback.threshold, len(self._names))
        safe_q = [(q or "") for q in queries]
        r = _cabi.host_features(safe_q, [len(c) if c else 0 for c in contexts], hp)
        ccat, scat = list(self.COMPLEX), list(self.SIMPLE)
        feats, toks = [], []
        for i in range(n):
            rule = int(r["rule"][i])
            if rule < 0:
                feats.append(self.classify(queries[i], contexts[i]))
                toks.append(self._token_fallback.count(queries[i], contexts[i]))
                continue
            dev, conf, tok = int(r["heur_dev"][i]), float(r["heur_conf"][i]), int(r["tok"][i])
            q = safe_q[i].strip()
            method, score = "heuristic", None
            if rule == 0:
                why = f"complex pattern={ccat[int(r['cat'][i])]}"
            elif rule == 1:
                why = f"long query chars={len(q)}"
            elif rule == 2:
                why = f"multi-question count={int(r['cat'][i])}"
            elif rule == 3:
                why = "code/debug markers detected"
            elif rule == 4:
                why = f"large context chars={len(contexts[i])}"
            elif rule == 5:
                why = f"simple pattern={scat[int(r['cat'][i])]}"
            elif rule == 6:
                why = "short everyday query"
            else:
                method, score = "heuristic_fallback", float(tok)
                why = f"no heuristic match -> tokens={tok} threshold={self._token_fallback.threshold}"
            feats.append((dev, conf, method, why, score))
            toks.append(tok)
        return feats, toks


# ----------------------------------------------------------------------------- hybrid
class HybridRouter(BaseRouter):
    """Weighted vote of token / semantic / heuristic (reference qre:371-414), evaluated in the
    GPU epilogue: S[dev] = sum_r w_r * conf_r * [dev_r == dev] (+ w_perf * H[dev]), argmax."""

    def __init__(self, config: Dict[str, Any], engine: Optional[Engine] = None):
        super().__init__(config)
        self.weights = config.get("weights", {"token": 0.35, "semantic": 0.35, "heuristic": 0.30})
        self._names = _device_names(config)
        self.routers = {
            "token": TokenBasedRouter(config),
            # the reference keys this on the sentence-transformers import (qre:380); an embedder supplied
            # through config["embedder"] serves as well
            "semantic": (SemanticRouter(config, engine=engine)
                         if SENTENCE_TRANSFORMERS_AVAILABLE or config.get("embedder") is not None else None),
            "heuristic": HeuristicRouter(config),
        }

    # host features of one query: everything the device pass needs besides the embedding
    def features(self, query: str, context: Optional[str]):
        tok_d = self.routers["token"].route(query, context)
        heur = self.routers["heuristic"].classify(query, context)
        return tok_d, heur

    def _decision(self, out: dict, i: int, tok_d: RoutingDecision, heur) -> RoutingDecision:
        sem = self.routers["semantic"]
        sem_d = sem._decision(out, i, tok_d)
        parts = [("token", tok_d.device, tok_d.confidence), ("semantic", sem_d.device, sem_d.confidence),
                 ("heuristic", self._names[heur[0]], heur[1])]
        S = out["score"][i]
        desc = " | ".join(f"{n}:{dev} conf={c:.2f} w={float(self.weights.get(n, 0.0)):.2f}" for n, dev, c in parts)
        return RoutingDecision(
            device=self._names[int(out["dev"][i])], confidence=float(out["conf"][i]), method="hybrid",
            reasoning=f"nano_score={float(S[0]):.3f} orin_score={float(S[-1]):.3f} | {desc}")

    def _host_vote(self, votes: Sequence[Tuple[str, int, float]]) -> RoutingDecision:
        """No embedder: only token + heuristic vote; tiny, stays on the host (qre:389-414)."""
        S = [0.0] * len(self._names)
        parts = []
        for name, dev, conf in votes:
            w = float(self.weights.get(name, 0.0))
            S[dev] += w * float(conf)
            parts.append(f"{name}:{self._names[dev]} conf={conf:.2f} w={w:.2f}")
        top = max(range(len(S)), key=lambda d: (S[d], -d))
        rest = [S[d] for d in range(len(S)) if d != top]
        margin = S[top] - (max(rest) if rest else 0.0)
        total = sum(S)
        conf = float(min(max(margin / total if total > 1e-12 else 0.5, 0.0), 1.0))
        return RoutingDecision(device=self._names[top], confidence=conf, method="hybrid",
                               reasoning=f"nano_score={S[0]:.3f} orin_score={S[-1]:.3f} | " + " | ".join(parts))

    def route(self, query: str, context: Optional[str] = None, q_emb: Optional[np.ndarray] = None
              ) -> RoutingDecision:
        tok_d, heur = self.features(query, context)
        sem = self.routers["semantic"]
        if sem is None:
            return self._host_vote([("token", self._names.index(tok_d.device), tok_d.confidence),
                                    ("heuristic", heur[0], heur[1])])
        if q_emb is None:
            q_emb = sem.embedder.encode([query])[0]
        out = sem._device_pass(q_emb, int(tok_d.complexity_score), (heur[0], heur[1]))
        return self._decision(out, 0, tok_d, heur)


# ----------------------------------------------------------------------------- perf
class PerformanceAwareRouter(BaseRouter):
    """Sliding-window latency/failure score per device, lower wins (reference qre:421-458)."""

    def __init__(self, config: Dict[str, Any]):
        super().__init__(config)
        self.window = int(config.get("perf_window", 30))
        self.fail_penalty = float(config.get("perf_fail_penalty", 3000.0))
        self._names = _device_names(config)
        self.stats: Dict[str, deque] = {n: deque(maxlen=self.window) for n in self._names}

    def update(self, device: str, latency_ms: float, tokens: int, ok: bool = True) -> None:
        if device in self.stats:
            self.stats[device].append((float(latency_ms), int(tokens), 1 if ok else 0))

    def _score(self, device: str) -> float:
        data = list(self.stats[device])
        if not data:
            return float("inf")
        lat = sum(x[0] for x in data)
        tok = sum(x[1] for x in data)
        fail_rate = 1.0 - (sum(x[2] for x in data) / len(data))
        base = lat / len(data) if tok == 0 else lat / tok
        return float(base + self.fail_penalty * fail_rate)

    def health(self) -> List[float]:
        """H[dev] in [0, 1] for the generalised hybrid score: 1 / (1 + score), 0 when unknown."""
        out = []
        for n in self._names:
            s = self._score(n)
            out.append(0.0 if s == float("inf") else 1.0 / (1.0 + max(s, 0.0)))
        return out

    def route(self, query: str, context: Optional[str] = None) -> RoutingDecision:
        scores = [self._score(n) for n in self._names]
        if all(s == float("inf") for s in scores):
            return RoutingDecision(device=self._names[0], confidence=0.2, method="perf",
                                   reasoning="no perf stats yet -> default nano")
        best = 0
        for d in range(1, len(scores)):
            if scores[d] < scores[best]:
                best = d
        device = self._names[best]
        return RoutingDecision(device=device, confidence=0.70, method="perf",
                               reasoning=f"scores nano={scores[0]:.2f} orin={scores[-1]:.2f} -> {device}")


# ----------------------------------------------------------------------------- QueryRouter
class QueryRouter:
    """Strategy selector over the B200 routing cache (reference qre:465-691): embed once,
    one fused device pass (cache lookup + routing scores), then the reference's control flow:
    confident hit -> cached prediction; heavy context over a cached light device, or a weak
    prediction -> re-route and record; miss -> route and insert."""

    AVAILABLE_STRATEGIES = {
        "token": TokenBasedRouter, "semantic": SemanticRouter, "heuristic": HeuristicRouter,
        "hybrid": HybridRouter, "perf": PerformanceAwareRouter,
    }

    def __init__(self, strategy: str = "token", config: Optional[Dict[str, Any]] = None):
        self.config = config or self._default_config()
        if strategy not in self.AVAILABLE_STRATEGIES:
            raise ValueError(f"Unknown strategy={strategy}. Available={list(self.AVAILABLE_STRATEGIES)}")
        self.strategy_name = strategy
        self.cache_enabled = bool(self.config.get("cache_enabled", True))
        max_size = int(self.config.get("cache_max_size", 500))
        # config["gpus"] = [0, 1, ...]: the cache is row-sharded over those GPUs (one process drives
        # them all); config["gpu"] = i or nothing: one GPU.  Additive keys, reference defaults otherwise.
        self._engine = _new_engine(self.config, max(max_size, 1), int(self.config.get("max_batch", 64)))
        self.router = self._make_router(strategy)
        self._cache = QueryCache(
            max_size=max_size,
            ttl_seconds=int(self.config.get("cache_ttl_seconds", 3600)),
            similarity_threshold=float(self.config.get("cache_similarity_threshold", 0.85)),
            use_semantic=bool(self.config.get("use_semantic_cache", True)),
            prediction_confidence_threshold=float(self.config.get("prediction_confidence_threshold", 0.70)),
            engine=self._engine,
        )
        self.cache_embedder = None
        if self.config.get("use_semantic_cache", True):
            self.cache_embedder = _make_embedder(self.config)

    def _make_router(self, strategy: str) -> BaseRouter:
        cls = self.AVAILABLE_STRATEGIES[strategy]
        if cls in (SemanticRouter, HybridRouter):
            return cls(self.config, engine=self._engine)
        return cls(self.config)

    @property
    def strategy(self) -> str:
        return self.strategy_name

    def _default_config(self) -> Dict[str, Any]:
        return {**BENCHMARK_CFG, "cache_ttl_seconds": 3600, "cache_max_size": 500,
                "cache_similarity_threshold": 0.85, "use_semantic_cache": True,
                "prediction_confidence_threshold": 0.70}

    # ------------------------------------------------------------------ core routing
    def _fusable(self) -> bool:
        r = self.router
        return isinstance(r, HybridRouter) and r.routers["semantic"] is not None or isinstance(r, SemanticRouter)

    def route_query(self, query: str, context: Optional[str] = None,
                    context_key: Optional[str] = None) -> RoutingDecision:
        ctxk = context_key or "default"
        q_emb: Optional[np.ndarray] = None
        if self.cache_enabled and self.cache_embedder is not None:
            try:
                q_emb = self.cache_embedder.encode([query])[0]
            except Exception as exc:
                logger.warning("Cache embedding failed, continuing without: %s", exc)

        routed: Optional[RoutingDecision] = None      # decision from the fused device pass

        def route_now() -> RoutingDecision:
            return routed if routed is not None else self._route(query, context, q_emb)

        if self.cache_enabled:
            if q_emb is not None and self._fusable():
                result, routed = self._fused_pass(query, context, ctxk, q_emb)
            else:
                result = self._cache.lookup(query, ctxk, q_emb)
            if result is not None:
                context_len = len(context) if context else 0
                threshold = int(self.config.get("heuristic_context_chars", 800))
                heavy_ctx = context_len >= threshold and result.predicted_device == DEVICE_NAMES[0]
                if heavy_ctx or result.use_hybrid_fallback:
                    reason = (f"context_len={context_len}>={threshold} overrides cached nano" if heavy_ctx
                              else f"low prediction confidence={result.predicted_confidence:.2f}")
                    decision = route_now()
                    self._cache.insert(query, ctxk, device=decision.device, confidence=decision.confidence,
                                       method=decision.method, q_emb=q_emb)
                    decision.reasoning = f"cache hit (hybrid re-route: {reason}) | " + decision.reasoning
                    decision.cache_hit = True
                    return decision
                age = int((datetime.now() - result.entry.timestamp).total_seconds())
                return RoutingDecision(
                    device=result.predicted_device, confidence=result.predicted_confidence,
                    method=f"{self.strategy_name}_cached",
                    reasoning=(f"cache hit age={age}s hits={result.entry.hit_count} "
                               f"predicted={result.predicted_device} conf={result.predicted_confidence:.2f} "
                               f"context_len={context_len} history={len(result.entry.routing_history)}"),
                    cache_hit=True)

        decision = route_now()
        if self.cache_enabled:
            self._cache.insert(query, ctxk, device=decision.device, confidence=decision.confidence,
                               method=decision.method, q_emb=q_emb)
        return decision

    def route_queries(self, queries: Sequence[str], contexts: Optional[Sequence[Optional[str]]] = None,
                      context_keys: Optional[Sequence[Optional[str]]] = None,
                      q_embs: Optional[np.ndarray] = None) -> List[RoutingDecision]:
        """Additive batch entry point (the reference routes one query per call, qre:559): every
        query of the batch is looked up against ONE snapshot of the cache and routed in a single
        fused device pass; the hit / override / insert control flow of `route_query` is then
        applied per query, in order, with all inserts after all lookups.  Equal to a loop of
        `route_query` whenever no query of the batch would hit an entry inserted by an earlier
        query of the same batch."""
        n = len(queries)
        contexts = list(contexts) if contexts is not None else [None] * n
        ctxks = [(k or "default") for k in (context_keys if context_keys is not None else [None] * n)]
        if not (self.cache_enabled and self._fusable()) or (q_embs is None and self.cache_embedder is None):
            return [self.route_query(q, c, k) for q, c, k in zip(queries, contexts, ctxks)]
        if q_embs is None:
            q_embs = np.asarray(self.cache_embedder.encode(list(queries)), np.float32)
        r = self.router
        hybrid = isinstance(r, HybridRouter)
        sem = r.routers["semantic"] if hybrid else r
        keys = None
        if hybrid:        # heuristic rules + token estimate for the whole batch on C++ threads, as arrays
            tok, hd, hc, keys = r.routers["heuristic"].features_batch(list(queries), contexts, ctxks)
            feats = (tok.tolist(), hd.tolist(), hc.tolist())
            route_in = dict(tok=tok, heur_dev=hd, heur_conf=hc)
        else:
            feats = [(r._token_fallback.route(q, c), (0, 0.0, "", "", None)) for q, c in zip(queries, contexts)]
            route_in = dict(tok=np.array([int(t.complexity_score) for t, _ in feats], np.int32),
                            heur_dev=np.array([h[0] for _, h in feats], np.uint8),
                            heur_conf=np.array([h[1] for _, h in feats], np.float64))
        with _gc_paused():
            results, out = self._cache._batch_pass(list(queries), ctxks, q_embs, route_in, sem._route_cfg(), keys,
                                                   as_tuples=True)
            with self._cache._deferred():        # all inserts of the batch reach the device in a few calls
                return self._finish_batch(queries, contexts, ctxks, q_embs, feats, results, out, hybrid)

    def _finish_batch(self, queries, contexts, ctxks, q_embs, feats, results, out, hybrid) -> List[RoutingDecision]:
        r = self.router
        decisions: List[RoutingDecision] = []
        threshold = int(self.config.get("heuristic_context_chars", 800))
        keys, own_emb = out.get("_keys"), out.get("_emb")
        now = datetime.now()
        cached_method = f"{self.strategy_name}_cached"
        if hybrid:
            # decisions straight from the result arrays; the reasoning string (and the token /
            # semantic sub-decisions it quotes) is only formatted if somebody reads it
            toks, hds, hcs = feats
            names, tr = r._names, r.routers["token"]
            devs, confs = out["dev"].tolist(), out["conf"].tolist()

            def why(i):
                tok_d = tr._decision(toks[i], *tr.decide(toks[i]))
                return r._decision(out, i, tok_d, (hds[i], hcs[i])).reasoning
        if hybrid and not any(results):
            # nothing was cached: every query takes the device's decision and is inserted
            new_obj, cls = object.__new__, LazyRoutingDecision
            dev_names = [names[d] for d in devs]
            for i in range(len(dev_names)):
                d = new_obj(cls)
                d.__dict__ = {"device": dev_names[i], "confidence": confs[i], "method": "hybrid",
                              "complexity_score": None, "cache_hit": False, "_lazy": (why, i)}
                decisions.append(d)
            self._cache.insert_batch(list(queries), ctxks, dev_names, confs, "hybrid", q_embs, _keys=keys,
                                     _emb_owned=own_emb)
            return decisions
        ins: List[int] = []                  # batch indices to insert, in order (all lookups came first)
        new_obj, lazy_cls, nano = object.__new__, LazyRoutingDecision, DEVICE_NAMES[0]
        for i, c in enumerate(contexts):
            res = results[i]                 # None, or (entry, predicted device, its confidence, weak prediction)
            routed = None
            if res is not None:
                e, pdev, pconf, weak = res
                context_len = len(c) if c else 0
                heavy_ctx = context_len >= threshold and pdev == nano
                if not (heavy_ctx or weak):
                    # the figures the reasoning quotes are taken NOW (hit count / history may move on)
                    d = new_obj(lazy_cls)
                    d.__dict__ = {"device": pdev, "confidence": pconf, "method": cached_method,
                                  "complexity_score": None, "cache_hit": True,
                                  "_lazy": (_cached_why, (now, e.timestamp, e.hit_count, pdev, pconf, context_len,
                                                          len(e.routing_history)))}
                    decisions.append(d)
                    continue
                reason = (f"context_len={context_len}>={threshold} overrides cached nano" if heavy_ctx
                          else f"low prediction confidence={pconf:.2f}")
                routed = self._routed(i, feats, out, hybrid)
                routed.reasoning = f"cache hit (hybrid re-route: {reason}) | " + routed.reasoning
                routed.cache_hit = True
            if routed is None:
                routed = (LazyRoutingDecision(names[devs[i]], confs[i], "hybrid", why, i)
                          if hybrid else self._routed(i, feats, out, hybrid))
            ins.append(i)
            decisions.append(routed)
        if ins:
            whole = len(ins) == len(decisions)
            pick = (lambda xs: xs) if whole else (lambda xs: [xs[i] for i in ins])
            routed_d = pick(decisions)
            methods = {d.method for d in routed_d}
            if len(methods) == 1:
                own = own_emb
                self._cache.insert_batch(pick(list(queries)), pick(ctxks), [d.device for d in routed_d],
                                         [d.confidence for d in routed_d], methods.pop(),
                                         q_embs if whole else q_embs[ins], _keys=pick(keys),
                                         _emb_owned=own if (whole or own is None) else own[ins])
            else:
                for i in ins:
                    d = decisions[i]
                    self._cache.insert(queries[i], ctxks[i], device=d.device, confidence=d.confidence,
                                       method=d.method, q_emb=q_embs[i], _key=keys[i])
        return decisions

    def _routed(self, i: int, feats, out, hybrid: bool) -> RoutingDecision:
        """The fully formatted routing decision of batch item i (the per-query path's form)."""
        r = self.router
        if not hybrid:
            return r._decision(out, i, feats[i][0])
        toks, hds, hcs = feats
        tr = r.routers["token"]
        return r._decision(out, i, tr._decision(toks[i], *tr.decide(toks[i])), (hds[i], hcs[i]))

    def _route(self, query: str, context: Optional[str], q_emb: Optional[np.ndarray]) -> RoutingDecision:
        if isinstance(self.router, HybridRouter) and q_emb is not None:
            return self.router.route(query, context, q_emb=q_emb)
        return self.router.route(query, context)

    def _fused_pass(self, query: str, context: Optional[str], ctxk: str, q_emb: np.ndarray):
        """ONE device pass: cache scan + centroid cosines + token/semantic/hybrid epilogue."""
        r = self.router
        if isinstance(r, HybridRouter):
            tok_d, heur = r.features(query, context)
            sem = r.routers["semantic"]
        else:
            tok_d, heur, sem = r._token_fallback.route(query, context), (0, 0.0, "", "", None), r
        route_in = dict(tok=np.array([int(tok_d.complexity_score)], np.int32),
                        heur_dev=np.array([heur[0]], np.uint8), heur_conf=np.array([heur[1]], np.float64))
        result, out = self._cache._lookup_and_route(query, ctxk, q_emb, route_in, sem._route_cfg())
        decision = r._decision(out, 0, tok_d, heur) if isinstance(r, HybridRouter) else r._decision(out, 0, tok_d)
        return result, decision

    # ------------------------------------------------------------------ cache passthroughs
    def warm_up_cache(self, pairs: List[Tuple[str, str, str]]) -> None:
        self._cache.warm_up(pairs, embedder=self.cache_embedder)

    def save_cache(self, path: str) -> None:
        self._cache.save(path)

    def load_cache(self, path: str) -> int:
        return self._cache.load(path)

    def invalidate_cache(self, context_key: Optional[str] = None, query_pattern: Optional[str] = None) -> int:
        return self._cache.invalidate(context_key=context_key, query_pattern=query_pattern)

    def get_cache_stats(self) -> Dict[str, Any]:
        return self._cache.stats()

    def clear_cache(self) -> None:
        self._cache.clear()

    # ------------------------------------------------------------------ perf + strategy
    def update_perf(self, device: str, latency_ms: float, tokens: int, ok: bool = True) -> None:
        if hasattr(self.router, "update"):
            self.router.update(device=device, latency_ms=latency_ms, tokens=tokens, ok=ok)

    def change_strategy(self, strategy: str) -> None:
        if strategy not in self.AVAILABLE_STRATEGIES:
            raise ValueError(f"Unknown strategy={strategy}")
        self.strategy_name = strategy
        self.router = self._make_router(strategy)


# Shared configs (reference qre:704-731): same keys and values.
BENCHMARK_CFG = dict(
    token_threshold=1000, model="meta-llama/Llama-2-7b-hf", embedding_model="all-MiniLM-L6-v2",
    semantic_label_path="src/tests/semantic_labels.json", semantic_margin_threshold=0.03,
    semantic_min_similarity=0.05, heuristic_long_chars=800, heuristic_multi_qmarks=2,
    heuristic_code_markers_needed=2, heuristic_context_chars=3200,
    weights=dict(token=0.25, semantic=0.45, heuristic=0.30), cache_enabled=False, perf_window=30,
    perf_fail_penalty=3000.0,
)

PRODUCTION_CFG = dict(
    BENCHMARK_CFG, cache_enabled=True, cache_ttl_seconds=3600, cache_max_size=500,
    cache_similarity_threshold=0.85, use_semantic_cache=True, prediction_confidence_threshold=0.70,
    enable_response_cache=True,
)
