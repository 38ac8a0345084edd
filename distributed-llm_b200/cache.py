"""cache.py — drop-in for the reference's `src/cache.py` with the store on a B200.

Same import surface (`from cache import CacheEntry, CacheLookupResult, QueryCache`,
reference query_router_engine.py:26) and the same observable behaviour; what changes is
where the work happens:

  * the embeddings live in a device-resident matrix (fp32 master rows + L2-normalised bf16
    rows, appended / refreshed / tombstoned in place) owned by `qre_b200.engine.Engine`;
  * the semantic scan of reference cache.py:267-305 is one call into libqre_b200.so
    (tcgen05 bf16 GEMM + fused top-k, exact fp32 re-score, see include/qre_b200.h);
  * the host keeps what is string- or history-shaped: the md5 exact-match index
    (cache.py:518-521), LRU order, TTL bookkeeping, routing history and its weighted vote
    (cache.py:94-140), regex invalidation, JSON persistence and stats.

Host-side complexity also drops: TTL expiry uses a heap instead of the reference's O(N) scan
per lookup (cache.py:250, 532-538), with identical eviction counts.

There is NO CPU fallback: without libqre_b200.so and an sm_100 GPU the first operation that
needs the device raises `qre_b200._cabi.QreError`.
"""
from __future__ import annotations

import gc
import hashlib
import heapq
import json
import logging
import re
import contextlib
import threading
from collections import OrderedDict
from dataclasses import dataclass, field
from datetime import datetime, timedelta
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

from qre_b200 import _cabi
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

PREDICTION_CONFIDENCE_THRESHOLD = 0.60      # reference cache.py:42
RECENCY_DECAY = 0.85                        # reference cache.py:46
DEVICE_NAMES: List[str] = ["nano", "orin"]  # index = device id on the GPU side

@contextlib.contextmanager
def _gc_paused():
    """The batch entry points create a few thousand small acyclic objects per call; a cyclic
    collection that happens to start in the middle walks the whole store (every entry, record and
    list of a large cache) and costs more than the batch.  Collection is switched back on — if it
    was on — when the batch is done."""
    was = gc.isenabled()
    gc.disable()
    try:
        yield
    finally:
        if was:
            gc.enable()


class _KeyBlock:
    """Payload of ONE expiry-heap record standing for many entries stamped at the same instant (a
    batch insert).  Never ordered before or after anything: heap ties between it and a plain key are
    left as they are."""
    __slots__ = ("keys",)

    def __init__(self, keys: List[str]):
        self.keys = keys

    def __lt__(self, other):
        return False

    def __gt__(self, other):
        return False


_EPOCH = datetime(1970, 1, 1)
_ONE_US = timedelta(microseconds=1)


def _to_us(t: datetime) -> int:
    """Exact microseconds since the epoch (naive datetimes, like the reference's now())."""
    return (t - _EPOCH) // _ONE_US


def device_index(name: str) -> int:
    """Anything that is not a known device counts as device 0, as the reference's vote treats
    every non-"orin" record as "nano" (cache.py:124-127)."""
    try:
        return DEVICE_NAMES.index(name)
    except ValueError:
        return 0


@dataclass
class RoutingRecord:
    """One routing decision remembered for a cached query (reference cache.py:53-76)."""
    device: str
    confidence: float
    method: str
    timestamp: str

    def to_dict(self) -> Dict[str, Any]:
        return dict(device=self.device, confidence=self.confidence, method=self.method,
                    timestamp=self.timestamp)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "RoutingRecord":
        return cls(d["device"], float(d["confidence"]), d.get("method", "unknown"),
                   d.get("timestamp", datetime.now().isoformat()))


@dataclass
class CacheEntry:
    """Same fields as the reference (cache.py:79-92); `_slot`/`_seq`/`_ts_us` are the row's
    place in the device matrix, its LRU sequence number and its timestamp in microseconds."""
    query: str
    query_hash: str
    context_key: str
    embedding: Optional[np.ndarray]
    timestamp: datetime
    device_used: str
    response_time: Optional[float] = None
    hit_count: int = 0
    routing_history: List[RoutingRecord] = field(default_factory=list)
    _slot: Optional[int] = field(default=None, repr=False, compare=False)
    _seq: int = field(default=0, repr=False, compare=False)
    _pred: Optional[tuple] = field(default=None, repr=False, compare=False)   # memo: (len, id of last, result)

    MAX_HISTORY = 20

    def record_routing(self, device: str, confidence: float, method: str) -> None:
        self.routing_history.append(RoutingRecord(device, confidence, method, datetime.now().isoformat()))
        del self.routing_history[:-self.MAX_HISTORY]
        self.device_used = device
        self._pred = None

    def predict_device(self) -> Tuple[str, float]:
        """Recency-weighted vote (reference cache.py:106-140): newest record weighs
        confidence, the i-th newest 0.85**i * confidence; the winner's share of the total is
        the confidence; equal scores go to the HIGHER device index ("orin" >= "nano");
        no history or ~zero weight -> (device_used, 0.5).  The vote only depends on the history,
        so it is memoised until the history (or device_used) changes."""
        h = self.routing_history
        memo = self._pred
        if memo is not None and memo[0] == len(h) and (not h or memo[1] is h[-1]) and memo[2] == self.device_used:
            return memo[3]
        res = self._predict_device()
        self._pred = (len(h), h[-1] if h else None, self.device_used, res)
        return res

    def _predict_device(self) -> Tuple[str, float]:
        if not self.routing_history:
            return self.device_used, 0.5
        score = [0.0] * len(DEVICE_NAMES)
        total = 0.0
        for age, rec in enumerate(reversed(self.routing_history)):
            w = (RECENCY_DECAY ** age) * rec.confidence
            score[device_index(rec.device)] += w
            total += w
        if total < 1e-9:
            return self.device_used, 0.5
        best = 0
        for d in range(1, len(score)):
            if score[d] >= score[best]:
                best = d
        return DEVICE_NAMES[best], float(min(score[best] / total, 1.0))

    def to_dict(self) -> Dict[str, Any]:
        return {
            "query": self.query, "query_hash": self.query_hash, "context_key": self.context_key,
            "embedding": None if self.embedding is None else self.embedding.tolist(),
            "timestamp": self.timestamp.isoformat(), "device_used": self.device_used,
            "response_time": self.response_time, "hit_count": self.hit_count,
            "routing_history": [r.to_dict() for r in self.routing_history],
        }

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "CacheEntry":
        emb = np.array(d["embedding"], dtype=np.float32) if d.get("embedding") else None
        return cls(query=d["query"], query_hash=d["query_hash"], context_key=d["context_key"],
                   embedding=emb, timestamp=datetime.fromisoformat(d["timestamp"]),
                   device_used=d["device_used"], response_time=d.get("response_time"),
                   hit_count=d.get("hit_count", 0),
                   routing_history=[RoutingRecord.from_dict(r) for r in d.get("routing_history", [])])


@dataclass
class CacheLookupResult:
    """What `lookup` returns on a hit (reference cache.py:179-192)."""
    entry: CacheEntry
    predicted_device: str
    predicted_confidence: float
    use_hybrid_fallback: bool


class QueryCache:
    """LRU + TTL routing cache with the semantic scan on the GPU.

    Constructor and method signatures are the reference's (cache.py:210-217, 236, 328, 372,
    398, 426, 434, 471, 502).  Keyword-only extras: `engine` (share a device store with the
    router so lookup and routing fuse into one launch sequence), `device` (GPU index),
    `max_batch` (largest `lookup_batch`), `gpus` (list of GPU indices: the rows are sharded over
    them and one process drives them all)."""

    def __init__(self, max_size: int = 100, ttl_seconds: int = 300, similarity_threshold: float = 0.85,
                 use_semantic: bool = True,
                 prediction_confidence_threshold: float = PREDICTION_CONFIDENCE_THRESHOLD, *,
                 engine: Optional[Any] = None, device: int = 0, max_batch: int = 64,
                 gpus: Optional[List[int]] = None):
        self.max_size = max_size
        self.ttl_seconds = ttl_seconds
        self.similarity_threshold = similarity_threshold
        self.use_semantic = use_semantic
        self.prediction_confidence_threshold = prediction_confidence_threshold

        self._store: "OrderedDict[str, CacheEntry]" = OrderedDict()
        self._lock = threading.RLock()
        self._engine = engine if engine is not None else _new_engine(
            {"gpus": gpus, "gpu": device}, max(max_size, 1), max_batch)
        self._slot_key: Dict[int, str] = {}          # device slot -> query_hash
        self._ctx_ids: Dict[str, int] = {}           # context_key -> u64 id the device compares
        self._seq = 0                                # LRU clock: larger = more recently used
        self._exp_heap: List[Tuple[int, str]] = []   # (ts_us, hash), lazy deletion
        self._stale_heap: List[Tuple[int, str, int]] = []   # (seq, hash, ts_us) of expired entries

        self._hits = 0
        self._attempts = 0
        self._evictions = 0
        self._hybrid_fallbacks = 0
        self._defer: Optional[dict] = None           # batched device sync in progress (see _deferred)

    # ------------------------------------------------------------------ batched device sync
    @contextlib.contextmanager
    def _deferred(self):
        """Inside this block the device mirror of host-side changes is queued and written at
        the end with a handful of batched C-ABI calls (one tombstone, one append, one or two
        updates) instead of one synchronous call per entry.  Only the batch entry points use it:
        they look everything up against a snapshot first, so nothing reads the device rows before
        the flush."""
        with self._lock:
            if self._defer is not None:              # nested: the outer block flushes
                yield
                return
            self._defer = {"dirty": OrderedDict(), "tomb": [], "lru": {}}
            try:
                yield
            finally:
                q, self._defer = self._defer, None
                self._flush(q)

    def _flush(self, q: dict) -> None:
        eng = self._engine
        if q["tomb"]:
            eng.tombstone(q["tomb"])
        new = [(e, rc) for e, rc in q["dirty"].values() if e._slot is None and e.embedding is not None]
        old = [(e, rc) for e, rc in q["dirty"].values() if e._slot is not None and e.embedding is not None]
        if new:
            pred = [e.predict_device() for e, _ in new]
            slots = eng.append_many(np.stack([e.embedding for e, _ in new]),
                                    [self._ctx_id(e.context_key) for e, _ in new],
                                    [_to_us(e.timestamp) for e, _ in new], [e._seq for e, _ in new],
                                    [device_index(d) for d, _ in pred], [c for _, c in pred])
            for (e, _), sl in zip(new, slots):
                e._slot = int(sl)
                self._slot_key[e._slot] = e.query_hash
        for changed in (True, False):
            grp = [e for e, rc in old if rc == changed]
            if not grp:
                continue
            pred = [e.predict_device() for e in grp]
            eng.update_many([e._slot for e in grp], rows=np.stack([e.embedding for e in grp]) if changed else None,
                            ts_us=[_to_us(e.timestamp) for e in grp], lru=[e._seq for e in grp],
                            pdev=[device_index(d) for d, _ in pred], pconf=[c for _, c in pred])
        touched = {e._slot for e, _ in q["dirty"].values() if e._slot is not None}
        lru = {s_: v for s_, v in q["lru"].items() if s_ not in touched and s_ in self._slot_key}
        if lru:
            eng.update_many(list(lru), lru=list(lru.values()))

    def _bump_lru(self, entry: CacheEntry) -> None:
        if entry._slot is None:
            return
        if self._defer is not None:
            self._defer["lru"][entry._slot] = entry._seq
        else:
            self._engine.update(entry._slot, lru=entry._seq)

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _make_hash(query: str, context_key: str) -> str:
        """md5 of "<ctx>||<query lowered and stripped>" (reference cache.py:518-521)."""
        return hashlib.md5(f"{context_key}||{query.lower().strip()}".encode("utf-8")).hexdigest()

    def _ttl_us(self) -> int:
        return int(self.ttl_seconds * 1_000_000)

    def _is_valid(self, entry: CacheEntry, now: Optional[datetime] = None) -> bool:
        now = datetime.now() if now is None else now
        return (now - entry.timestamp).total_seconds() <= self.ttl_seconds

    def _ctx_id(self, context_key: str) -> int:
        cid = self._ctx_ids.get(context_key)
        if cid is None:
            cid = len(self._ctx_ids) + 1             # never the tombstone value
            self._ctx_ids[context_key] = cid
        return cid

    def _cfg(self) -> _cabi.RouteConfig:
        return _cabi.RouteConfig(similarity_threshold=float(self.similarity_threshold), ttl_us=self._ttl_us())

    def _touch(self, key: str, entry: CacheEntry) -> None:
        """move_to_end + the matching lru_seq on the device row."""
        self._store.move_to_end(key)
        self._seq += 1
        entry._seq = self._seq

    def _stamp(self, key: str, entry: CacheEntry, now: datetime) -> None:
        entry.timestamp = now
        heapq.heappush(self._exp_heap, (_to_us(now), key))

    def _delete_entry(self, key: str) -> None:
        entry = self._store.pop(key, None)
        if entry is None:
            return
        if self._defer is not None:
            self._defer["dirty"].pop(key, None)
        if entry._slot is not None:
            self._slot_key.pop(entry._slot, None)
            if self._defer is not None:
                self._defer["tomb"].append(entry._slot)
                self._defer["lru"].pop(entry._slot, None)
            else:
                self._engine.tombstone([entry._slot])
            entry._slot = None
        entry.embedding = None

    def _collect_stale(self, now_us: int) -> None:
        limit = now_us - self._ttl_us()
        while self._exp_heap and self._exp_heap[0][0] < limit:
            ts, key = heapq.heappop(self._exp_heap)
            for k in (key.keys if type(key) is _KeyBlock else (key,)):
                e = self._store.get(k)
                if e is not None and _to_us(e.timestamp) == ts:
                    heapq.heappush(self._stale_heap, (e._seq, k, ts))

    def _pop_stale(self, now_us: int) -> Optional[str]:
        """Earliest-in-LRU-order entry that is expired right now, or None."""
        limit = now_us - self._ttl_us()
        while self._stale_heap:
            seq, key, ts = heapq.heappop(self._stale_heap)
            e = self._store.get(key)
            if e is not None and e._seq == seq and _to_us(e.timestamp) == ts and ts < limit:
                return key
        return None

    def _evict_expired(self, now: Optional[datetime] = None) -> None:
        """Drop every TTL-expired entry, counting each (reference cache.py:532-538)."""
        with self._lock:
            now_us = _to_us(datetime.now() if now is None else now)
            self._collect_stale(now_us)
            key = self._pop_stale(now_us)
            if key is None:
                return
            with self._deferred():                   # a wave of expiries frees its device rows in one call
                while key is not None:
                    self._delete_entry(key)
                    self._evictions += 1
                    key = self._pop_stale(now_us)

    def _evict_one(self, now: datetime) -> None:
        """At capacity: first expired entry in LRU order, else the LRU front
        (reference cache.py:540-554)."""
        now_us = _to_us(now)
        self._collect_stale(now_us)
        key = self._pop_stale(now_us)
        if key is None and self._store:
            key = next(iter(self._store))
        if key is not None:
            self._delete_entry(key)
            self._evictions += 1

    def _sync_row(self, entry: CacheEntry, row_changed: bool) -> None:
        """Mirror an entry's scan-relevant state into its device row."""
        if entry.embedding is None:
            return
        if self._defer is not None:
            prev = self._defer["dirty"].get(entry.query_hash)
            self._defer["dirty"][entry.query_hash] = (entry, row_changed or (prev is not None and prev[1]))
            return
        pdev, pconf = entry.predict_device()
        ts_us = _to_us(entry.timestamp)
        if entry._slot is None:
            entry._slot = self._engine.append(entry.embedding, self._ctx_id(entry.context_key), ts_us,
                                              entry._seq, device_index(pdev), pconf)
            self._slot_key[entry._slot] = entry.query_hash
        else:
            self._engine.update(entry._slot, row=entry.embedding if row_changed else None, ts_us=ts_us,
                                lru=entry._seq, pdev=device_index(pdev), pconf=pconf)

    # ------------------------------------------------------------------ public API
    def lookup(self, query: str, context_key: str,
               q_emb: Optional[np.ndarray] = None) -> Optional[CacheLookupResult]:
        """Exact-hash match first, then the semantic scan (reference cache.py:236-326).
        Returns None on a miss."""
        return self._lookup_and_route(query, context_key, q_emb, None)[0]

    def _lookup_and_route(self, query: str, context_key: str, q_emb: Optional[np.ndarray],
                          route_in: Optional[dict], route_cfg: Optional[_cabi.RouteConfig] = None):
        """lookup() plus, when `route_in` is given, the routing epilogue in the same device
        pass.  Returns (CacheLookupResult | None, device outputs | None)."""
        with self._lock:
            self._attempts += 1
            now = datetime.now()
            self._evict_expired(now)
            key = self._make_hash(query, context_key)
            entry: Optional[CacheEntry] = None
            exact = self._store.get(key)
            if exact is not None and exact.context_key == context_key:
                if self._is_valid(exact, now):
                    entry = exact
                else:
                    self._delete_entry(key)
            want_scan = entry is None and self.use_semantic and q_emb is not None
            dev_out = None
            if q_emb is not None and (want_scan or route_in is not None):
                cfg = route_cfg if route_cfg is not None else self._cfg()
                cfg.similarity_threshold = float(self.similarity_threshold)
                cfg.ttl_us = self._ttl_us()
                dev_out = self._engine.scan(q_emb, np.array([self._ctx_id(context_key)], np.uint64),
                                            _to_us(now), cfg, route_in)
            if entry is None:
                if not want_scan:
                    return None, dev_out
                slot = int(dev_out["hit_slot"][0])
                if slot < 0:
                    return None, dev_out
                found = self._store.get(self._slot_key.get(slot, ""))
                if found is None or not self._is_valid(found, now):
                    return None, dev_out
                entry, key = found, found.query_hash
            entry.hit_count += 1
            self._touch(key, entry)
            self._bump_lru(entry)
            self._hits += 1
            pdev, pconf = entry.predict_device()
            fallback = pconf < self.prediction_confidence_threshold
            if fallback:
                self._hybrid_fallbacks += 1
                logger.debug("cache hit with weak prediction %.2f for %r", pconf, query[:60])
            return CacheLookupResult(entry, pdev, pconf, fallback), dev_out

    def lookup_batch(self, queries: List[str], context_keys: List[str],
                     q_embs: np.ndarray) -> List[Optional[CacheLookupResult]]:
        """Additive batch form: every query is scored against ONE snapshot of the cache (a
        single device pass), then hits are applied in order.  Equals a loop of `lookup` unless
        an earlier hit's LRU promotion would change a later tie-break."""
        return self._batch_pass(queries, context_keys, q_embs, None, None)[0]

    def _batch_pass(self, queries: List[str], context_keys: List[str], q_embs: np.ndarray,
                    route_in: Optional[dict], route_cfg: Optional[_cabi.RouteConfig],
                    keys: Optional[List[Optional[str]]] = None, as_tuples: bool = False):
        """Snapshot lookup of a whole batch (+ routing epilogue when `route_in` is given) in one
        device pass per `max_batch` chunk.  Returns (results, device outputs)."""
        with self._lock:
            now = datetime.now()
            self._evict_expired(now)
            q_embs = np.ascontiguousarray(q_embs, np.float32)
            ctx = np.array([self._ctx_id(c) for c in context_keys], np.uint64)
            cfg = route_cfg if route_cfg is not None else self._cfg()
            cfg.similarity_threshold = float(self.similarity_threshold)
            cfg.ttl_us = self._ttl_us()
            step = max(1, int(getattr(self._engine, "max_batch", 64)))

            def host_side():
                # what does not need the device's answer: the exact-match keys (one md5 per query,
                # unless the C++ feature pass handed them in) and the cache's own copy of the
                # embedding block, which a later insert_batch of this batch keeps
                nonlocal keys
                if keys is None:
                    keys = [self._make_hash(q, c) for q, c in zip(queries, context_keys)]
                else:
                    keys = [k if k is not None else self._make_hash(q, c)
                            for k, q, c in zip(keys, queries, context_keys)]
                side["emb"] = np.array(q_embs, dtype=np.float32, copy=True) if route_in is not None else None

            side: Dict[str, Any] = {"emb": None}
            overlapped = getattr(self._engine, "scan_overlapped", None)
            if overlapped is not None and 0 < len(queries) <= step:
                out = overlapped(q_embs, ctx, _to_us(now), cfg, route_in, host_side)    # host_side runs under the scan
            else:
                outs = []
                for i0 in range(0, len(queries), step):
                    sl = slice(i0, min(len(queries), i0 + step))
                    ri = None if route_in is None else {k: v[sl] for k, v in route_in.items()}
                    outs.append(self._engine.scan(q_embs[sl], ctx[sl], _to_us(now), cfg, ri))
                out = {k: np.concatenate([o[k] for o in outs]) for k in outs[0]} if outs else {}
                host_side()
            results: List[Optional[CacheLookupResult]] = []
            with self._deferred():
                self._apply_batch_hits(queries, context_keys, keys, out, now, results, as_tuples)
            # the batch's keys and the private embedding copy travel with the device outputs (by value:
            # another thread's batch may run before the caller has finished with this one)
            out["_keys"], out["_emb"] = keys, side["emb"]
            return results, out

    def _apply_batch_hits(self, queries, context_keys, keys, out, now, results, as_tuples=False) -> None:
        """Host bookkeeping of a batch's hits, in order (exact-hash match first, then the device
        scan's row), as `lookup` does for one query."""
        slots = out["hit_slot"].tolist() if self.use_semantic else [-1] * len(queries)
        store, slot_key = self._store, self._slot_key
        ttl = self.ttl_seconds
        thr = self.prediction_confidence_threshold
        lru = self._defer["lru"] if self._defer is not None else None
        move, append, new_obj = store.move_to_end, results.append, object.__new__
        self._attempts += len(queries)
        for i, c in enumerate(context_keys):
            entry = store.get(keys[i])
            if entry is None or entry.context_key != c or (now - entry.timestamp).total_seconds() > ttl:
                entry = None
                slot = slots[i]
                if slot >= 0:
                    entry = store.get(slot_key.get(slot, ""))
                    if entry is not None and (now - entry.timestamp).total_seconds() > ttl:
                        entry = None
            if entry is None:
                append(None)
                continue
            entry.hit_count += 1
            move(entry.query_hash)                     # _touch: LRU order on the host ...
            self._seq = entry._seq = self._seq + 1
            if entry._slot is not None:                # ... and the matching lru_seq on the device row
                if lru is not None:
                    lru[entry._slot] = entry._seq
                else:
                    self._engine.update(entry._slot, lru=entry._seq)
            self._hits += 1
            memo = entry._pred                           # predict_device(), memo check inlined
            h = entry.routing_history
            if memo is not None and memo[0] == len(h) and (not h or memo[1] is h[-1]) and memo[2] == entry.device_used:
                pdev, pconf = memo[3]
            else:
                pdev, pconf = entry.predict_device()
            fb = pconf < thr
            if fb:
                self._hybrid_fallbacks += 1
            if as_tuples:                              # the router's batch path unpacks it at once
                append((entry, pdev, pconf, fb))
                continue
            r = new_obj(CacheLookupResult)
            r.__dict__ = {"entry": entry, "predicted_device": pdev, "predicted_confidence": pconf,
                          "use_hybrid_fallback": fb}
            append(r)

    def insert(self, query: str, context_key: str, device: str, confidence: float = 1.0,
               method: str = "unknown", q_emb: Optional[np.ndarray] = None,
               response_time: Optional[float] = None, *, _key: Optional[str] = None) -> None:
        """Insert, or refresh in place when the hash exists (reference cache.py:328-370).
        (`_key`: the batch path hands over the md5 it already computed for the lookup.)"""
        key = _key if _key is not None else self._make_hash(query, context_key)
        with self._lock:
            now = datetime.now()
            entry = self._store.get(key)
            if entry is not None:
                self._stamp(key, entry, now)
                entry.record_routing(device, confidence, method)
                if q_emb is not None:
                    entry.embedding = np.array(q_emb, dtype=np.float32, copy=True)
                if response_time is not None:
                    entry.response_time = response_time
                self._touch(key, entry)
                self._sync_row(entry, row_changed=q_emb is not None)
                return
            if len(self._store) >= self.max_size:
                self._evict_one(now)
            entry = CacheEntry(query=query, query_hash=key, context_key=context_key,
                               embedding=None if q_emb is None else np.array(q_emb, dtype=np.float32, copy=True),
                               timestamp=now, device_used=device, response_time=response_time)
            entry.record_routing(device, confidence, method)
            self._store[key] = entry
            self._stamp(key, entry, now)
            self._touch(key, entry)
            self._sync_row(entry, row_changed=True)

    def insert_batch(self, queries: Sequence[str], context_keys: Sequence[str], devices: Sequence[str],
                     confidences: Sequence[float], method: str = "unknown",
                     q_embs: Optional[np.ndarray] = None, *, _keys: Optional[List[str]] = None,
                     _emb_owned: Optional[np.ndarray] = None) -> None:
        """Additive batch form of `insert` (the reference inserts one entry per call,
        cache.py:328): the same entries, in the same order, as a loop of `insert`, with the host
        bookkeeping of brand-new entries done without per-entry device calls or per-entry array
        copies — one copy of the embedding block (entries hold row views of it), one timestamp for
        the whole batch, one `append` to the device for all new rows.  Keys that already exist,
        entries that need an eviction first and NaN confidences take `insert` itself."""
        n = len(queries)
        if n == 0:
            return
        keys = _keys if _keys is not None else [self._make_hash(q, c) for q, c in zip(queries, context_keys)]
        with self._lock, self._deferred():
            if (_emb_owned is not None and _emb_owned.dtype == np.float32 and _emb_owned.ndim == 2
                    and _emb_owned.shape[0] == n and _emb_owned.flags.c_contiguous):
                emb = _emb_owned                     # a private copy made earlier (under the device pass)
            else:
                emb = None if q_embs is None else np.array(q_embs, dtype=np.float32, copy=True)
            if emb is None or emb.ndim != 2:
                for i in range(n):
                    self.insert(queries[i], context_keys[i], devices[i], confidences[i], method,
                                None if emb is None else emb[i], _key=keys[i])
                return
            now = datetime.now()
            ts_us, iso = _to_us(now), now.isoformat()
            store = self._store
            new_obj = object.__new__
            ent: List[CacheEntry] = []               # entries created here, in order
            pos: Optional[List[int]] = None          # their batch indices, once some item went the slow way
            room = self.max_size - len(store)
            seq = self._seq
            for i in range(n):
                key, conf, dev = keys[i], confidences[i], devices[i]
                if key in store or room <= 0 or conf != conf:
                    if pos is None:
                        pos = list(range(i))
                    self._seq = seq
                    self.insert(queries[i], context_keys[i], dev, conf, method, emb[i], _key=key)
                    seq = self._seq
                    room = self.max_size - len(store)
                    continue
                room -= 1
                seq += 1
                rec = new_obj(RoutingRecord)
                rec.__dict__ = {"device": dev, "confidence": conf, "method": method, "timestamp": iso}
                # (response_time, hit_count, _slot keep their class defaults; with one record the
                # recency vote is (device, 1.0), or (device, 0.5) without weight)
                e = new_obj(CacheEntry)
                e.__dict__ = {"query": queries[i], "query_hash": key, "context_key": context_keys[i],
                              "embedding": emb[i], "timestamp": now, "device_used": dev, "routing_history": [rec],
                              "_seq": seq, "_pred": (1, rec, dev, (dev, 1.0 if conf >= 1e-9 else 0.5))}
                store[key] = e
                ent.append(e)
                if pos is not None:
                    pos.append(i)
            self._seq = seq
            if ent:                                  # one expiry record for the lot (same timestamp)
                heapq.heappush(self._exp_heap, (ts_us, _KeyBlock([e.query_hash for e in ent])))
            if pos is not None:
                # (an entry evicted again later in this same batch has lost its embedding: no row for it)
                keep = [j for j, e in enumerate(ent) if e.embedding is not None and e._slot is None]
                ent, pos = [ent[j] for j in keep], [pos[j] for j in keep]
            if not ent:
                return
            d = self._defer
            if d is not None and d["tomb"]:              # free the device slots of this batch's evictions first
                self._engine.tombstone(d["tomb"])
                d["tomb"] = []
            rows = emb if pos is None else emb[pos]
            ctx_of = {ck: self._ctx_id(ck) for ck in dict.fromkeys(e.context_key for e in ent)}   # first-seen order
            dev_of = {dv: device_index(dv) for dv in set(devices)}
            m = len(ent)
            if pos is None:
                ctx = [ctx_of[ck] for ck in context_keys]
                pdev = [dev_of[dv] for dv in devices]
                pconf = np.where(np.asarray(confidences, np.float64) >= 1e-9, 1.0, 0.5)
                lru = np.arange(seq - m + 1, seq + 1, dtype=np.uint64)
            else:
                pred = [e.predict_device() for e in ent]     # (a later item of the batch may have refreshed one)
                ctx = [ctx_of[e.context_key] for e in ent]
                pdev = [device_index(p[0]) for p in pred]
                pconf = [p[1] for p in pred]
                lru = [e._seq for e in ent]
            slots = self._engine.append_many(rows, ctx, np.full(m, ts_us, np.int64), lru, pdev, pconf)
            slot_key = self._slot_key
            for e, sl in zip(ent, slots.tolist() if hasattr(slots, "tolist") else slots):
                e._slot = sl
                slot_key[sl] = e.query_hash

    def invalidate(self, context_key: Optional[str] = None, query_pattern: Optional[str] = None) -> int:
        """Remove entries by context and/or case-insensitive regex on the query text; neither
        means everything (reference cache.py:372-396).  Returns the number removed."""
        rx = re.compile(query_pattern, re.IGNORECASE) if query_pattern else None
        with self._lock:
            doomed = [k for k, e in self._store.items()
                      if (context_key is None or e.context_key == context_key)
                      and (rx is None or rx.search(e.query))]
            with self._deferred():                   # one tombstone call for the lot
                for k in doomed:
                    self._delete_entry(k)
        logger.debug("invalidate removed %d entries", len(doomed))
        return len(doomed)

    def warm_up(self, pairs: List[Tuple[str, str, str]], embedder: Any = None) -> None:
        """Pre-populate from (query, context_key, device) triples (reference cache.py:398-420)."""
        vecs: List[Optional[np.ndarray]] = [None] * len(pairs)
        if embedder is not None:
            try:
                vecs = list(embedder.encode([q for q, _, _ in pairs]))
            except Exception as exc:
                logger.warning("warm_up: embedding failed, inserting without vectors: %s", exc)
        if pairs and all(v is not None for v in vecs):
            # the same entries in the same order as the loop below, with one device append
            self.insert_batch([q for q, _, _ in pairs], [c for _, c, _ in pairs], [d for _, _, d in pairs],
                              [1.0] * len(pairs), "unknown", np.asarray(vecs, dtype=np.float32))
        else:
            with self._deferred():
                for (q, ctx, dev), v in zip(pairs, vecs):
                    self.insert(q, ctx, dev, q_emb=v)
        logger.info("cache warmed up with %d entries", len(pairs))

    # ------------------------------------------------------------------ persistence
    def save(self, path: str) -> None:
        """JSON list of the non-expired entries, same schema as the reference
        (cache.py:426-432, 146-157), so files move between the two implementations."""
        self._evict_expired()
        with self._lock:
            data = [e.to_dict() for e in self._store.values()]
        Path(path).write_text(json.dumps(data, indent=2), encoding="utf-8")
        logger.info("cache saved %d entries -> %s", len(data), path)

    def load(self, path: str) -> int:
        """Restore entries; expired or malformed ones are skipped; returns the count loaded
        (reference cache.py:434-465; like the reference, max_size is not enforced here)."""
        p = Path(path)
        if not p.exists():
            logger.warning("cache load: %s not found", path)
            return 0
        try:
            raw = json.loads(p.read_text(encoding="utf-8"))
        except json.JSONDecodeError as exc:
            logger.error("cache load: bad JSON: %s", exc)
            return 0
        loaded = 0
        with self._lock, self._deferred():       # the rows reach the device in one append at the end
            now = datetime.now()
            for d in raw:
                try:
                    entry = CacheEntry.from_dict(d)
                except Exception as exc:
                    logger.warning("cache load: skipping malformed entry: %s", exc)
                    continue
                if not self._is_valid(entry, now):
                    continue
                old = self._store.get(entry.query_hash)
                if old is not None:                   # same key: replaced in place, position kept
                    entry._seq = old._seq
                    if old._slot is not None:
                        self._slot_key.pop(old._slot, None)
                        self._defer["tomb"].append(old._slot)
                        self._defer["lru"].pop(old._slot, None)
                        old._slot = None
                    self._store[entry.query_hash] = entry
                else:
                    self._store[entry.query_hash] = entry
                    self._seq += 1
                    entry._seq = self._seq
                heapq.heappush(self._exp_heap, (_to_us(entry.timestamp), entry.query_hash))
                self._ensure_capacity(len(self._store))
                self._sync_row(entry, row_changed=True)
                loaded += 1
        logger.info("cache loaded %d entries from %s", loaded, path)
        return loaded

    def save_snapshot(self, path: str) -> None:
        """Binary companion of `save` for large caches: the device rows and row
        metadata go to `<path>.rows` through qre_cache_snapshot_save (no float -> text), the
        host-side state (queries, hashes, histories, LRU order, slots) to `<path>` as JSON
        without embeddings."""
        self._evict_expired()
        with self._lock:
            if getattr(self._engine, "n", 1) > 1:
                raise _cabi.QreError("save_snapshot / load_snapshot work on a one-GPU cache; use save / load")
            if self._engine.shard is None:
                raise _cabi.QreError("nothing on the device to snapshot")
            self._engine.shard.snapshot_save(path + ".rows")
            entries = []
            for e in self._store.values():
                d = e.to_dict()
                d["embedding"] = None
                d["_slot"], d["_seq"] = e._slot, e._seq
                entries.append(d)
            meta = {"version": 1, "ctx_ids": self._ctx_ids, "seq": self._seq, "entries": entries,
                    "dim": self._engine.shard.dim}
        Path(path).write_text(json.dumps(meta), encoding="utf-8")

    def load_snapshot(self, path: str) -> int:
        """Restore a `save_snapshot` pair into an EMPTY cache; expired entries are dropped.
        Returns the number of entries loaded."""
        meta = json.loads(Path(path).read_text(encoding="utf-8"))
        with self._lock:
            if self._store:
                raise ValueError("load_snapshot needs an empty cache")
            eng = self._engine
            need = max([(d["_slot"] or 0) + 1 for d in meta["entries"]] + [1])
            if need > eng.capacity:
                eng.close()
                eng.capacity = need
            sh = eng.ensure(int(meta["dim"]))
            sh.clear()
            sh.snapshot_load(path + ".rows")
            self._ctx_ids = {k: int(v) for k, v in meta["ctx_ids"].items()}
            self._seq = int(meta["seq"])
            now = datetime.now()
            stale = []
            for d in meta["entries"]:
                entry = CacheEntry.from_dict(d)
                entry._slot, entry._seq = d["_slot"], int(d["_seq"])
                if not self._is_valid(entry, now):
                    if entry._slot is not None:
                        stale.append(entry._slot)
                    continue
                self._store[entry.query_hash] = entry
                if entry._slot is not None:
                    self._slot_key[entry._slot] = entry.query_hash
                heapq.heappush(self._exp_heap, (_to_us(entry.timestamp), entry.query_hash))
            if stale:
                eng.tombstone(stale)
            live = [e for e in self._store.values() if e._slot is not None]
            if live:
                rows = eng.read_rows([e._slot for e in live])
                for e, r in zip(live, rows):
                    e.embedding = r.copy()
            return len(self._store)

    def _ensure_capacity(self, need: int) -> None:
        """`load` may exceed max_size (reference quirk); rebuild the device store larger."""
        eng = self._engine
        if need <= eng.capacity:
            return
        new_cap = max(need, eng.capacity * 2)
        if eng.shard is not None:
            eng.close()
        eng.capacity = new_cap
        if self._defer is not None:                  # queued row operations named slots of the old store
            self._defer["tomb"].clear()
            self._defer["lru"].clear()
        self._slot_key.clear()
        for e in self._store.values():
            e._slot = None
        for e in self._store.values():
            self._sync_row(e, row_changed=True)

    # ------------------------------------------------------------------ stats
    def stats(self) -> Dict[str, Any]:
        """Same keys as the reference (cache.py:471-500)."""
        with self._lock:
            now = datetime.now()
            valid = sum(1 for e in self._store.values() if self._is_valid(e, now))
            hot = sorted(self._store.values(), key=lambda e: e.hit_count, reverse=True)[:5]
            size = len(self._store)
        return {
            "size": size, "max_size": self.max_size, "valid": valid, "stale": size - valid,
            "hits": self._hits, "attempts": self._attempts,
            "hit_rate": round(self._hits / max(self._attempts, 1), 4),
            "evictions": self._evictions, "hybrid_fallbacks": self._hybrid_fallbacks,
            "top_queries": [{"query": e.query[:60], "hits": e.hit_count,
                             "predicted_device": e.predict_device()[0],
                             "predicted_confidence": round(e.predict_device()[1], 3),
                             "history_len": len(e.routing_history)} for e in hot],
        }

    def clear(self) -> None:
        """Wipe the cache and reset the counters (reference cache.py:502-512)."""
        with self._lock:
            for e in self._store.values():
                e.embedding = None
                e._slot = None
            self._store.clear()
            self._slot_key.clear()
            self._exp_heap.clear()
            self._stale_heap.clear()
            self._engine.clear()
            self._hits = self._attempts = self._evictions = self._hybrid_fallbacks = 0
