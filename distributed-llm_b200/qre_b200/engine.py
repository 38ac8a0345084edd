"""Engine — the device side of one routing cache + its router state, behind the small store
interface the host classes (`cache.QueryCache`, `query_router_engine.SemanticRouter`) use.

The Shard is created lazily, once the embedding dimension is known (the reference learns it
from the first embedding it sees, src/cache.py:363).  There is no CPU fallback: creating the
Shard without an sm_100 GPU raises `QreError`.
"""
from __future__ import annotations

import functools
import threading
from typing import Dict, Optional, Sequence

import numpy as np

from . import _cabi


def _locked(fn):
    """A qre_cache_t handle is not thread-safe (include/qre_b200.h): the cache and the routers
    that share one Engine may be called from different request threads, so every native call
    goes through the engine's lock."""
    @functools.wraps(fn)
    def wrapper(self, *a, **k):
        with self._lock:
            return fn(self, *a, **k)
    return wrapper


class Engine:
    def __init__(self, capacity: int = 512, device: int = 0, max_batch: int = 64,
                 dim: Optional[int] = None, scan_mode: int = _cabi.SCAN_TENSOR):
        self.capacity = max(int(capacity), 1)
        self.device, self.max_batch, self.scan_mode = int(device), int(max_batch), int(scan_mode)
        self.shard: Optional[_cabi.Shard] = None
        self._centroids: Optional[np.ndarray] = None
        self._owner: Optional[np.ndarray] = None
        self._health: Optional[np.ndarray] = None
        self._lock = threading.RLock()
        if dim is not None:
            self.ensure(dim)

    # ---- lifetime ----
    @_locked
    def ensure(self, dim: int) -> _cabi.Shard:
        if self.shard is None:
            self.shard = _cabi.Shard(dim=int(dim), capacity=self.capacity, max_batch=self.max_batch,
                                     device=self.device)
            if self._centroids is not None:
                self.shard.set_centroids(self._centroids, self._owner)
            if self._health is not None:
                self.shard.set_perf_health(self._health)
        elif self.shard.dim != int(dim):
            raise ValueError(f"embedding dimension changed: cache holds {self.shard.dim}-d rows, got {dim}")
        return self.shard

    @_locked
    def close(self) -> None:
        if self.shard is not None:
            self.shard.close()
            self.shard = None

    # ---- router state ----
    @_locked
    def set_centroids(self, centroids: np.ndarray, owner: Sequence[int]) -> None:
        self._centroids = np.ascontiguousarray(centroids, np.float32)
        self._owner = np.ascontiguousarray(owner, np.uint8)
        if self.shard is None:
            self.ensure(self._centroids.shape[1])
        else:
            self.shard.set_centroids(self._centroids, self._owner)

    @_locked
    def set_perf_health(self, health: Sequence[float]) -> None:
        self._health = np.ascontiguousarray(health, np.float64)
        if self.shard is not None:
            self.shard.set_perf_health(self._health)

    @property
    def n_centroids(self) -> int:
        return 0 if self._centroids is None else int(self._centroids.shape[0])

    # ---- store interface used by QueryCache ----
    @_locked
    def append(self, row, ctx: int, ts_us: int, lru: int, pdev: int, pconf: float) -> int:
        row = np.ascontiguousarray(row, np.float32).reshape(1, -1)
        sh = self.ensure(row.shape[1])
        return int(sh.append(row, [ctx], [ts_us], [lru], [pdev], [pconf])[0])

    @_locked
    def append_many(self, rows, ctx, ts_us, lru, pdev, pconf) -> np.ndarray:
        rows = np.ascontiguousarray(rows, np.float32)
        return self.ensure(rows.shape[1]).append(rows, ctx, ts_us, lru, pdev, pconf)

    @_locked
    def update(self, slot: int, row=None, ts_us=None, lru=None, pdev=None, pconf=None) -> None:
        one = lambda v, dt: None if v is None else np.array([v], dt)   # noqa: E731
        self.shard.update(np.array([slot], np.int64),
                          rows=None if row is None else np.ascontiguousarray(row, np.float32).reshape(1, -1),
                          ts_us=one(ts_us, np.int64), lru_seq=one(lru, np.uint64),
                          pred_dev=one(pdev, np.uint8), pred_conf=one(pconf, np.float64))

    @_locked
    def update_many(self, slots, rows=None, ts_us=None, lru=None, pdev=None, pconf=None) -> None:
        arr = lambda v, dt: None if v is None else np.asarray(v, dt)   # noqa: E731
        self.shard.update(np.asarray(slots, np.int64),
                          rows=None if rows is None else np.ascontiguousarray(rows, np.float32),
                          ts_us=arr(ts_us, np.int64), lru_seq=arr(lru, np.uint64), pred_dev=arr(pdev, np.uint8),
                          pred_conf=arr(pconf, np.float64))

    @_locked
    def tombstone(self, slots) -> None:
        slots = np.asarray(list(slots), np.int64)
        if slots.size and self.shard is not None:
            self.shard.tombstone(slots)

    @_locked
    def clear(self) -> None:
        if self.shard is not None:
            self.shard.clear()

    @_locked
    def read_rows(self, slots) -> np.ndarray:
        return self.shard.read_rows(np.asarray(list(slots), np.int64))

    @_locked
    def scan(self, q, q_ctx, now_us: int, cfg: _cabi.RouteConfig, route_in: Optional[dict] = None) -> dict:
        """One fused device pass for a batch: lookup (always) + route (when route_in is given:
        dict(tok, heur_dev, heur_conf))."""
        q = np.ascontiguousarray(q, np.float32)
        if q.ndim == 1:
            q = q.reshape(1, -1)
        sh = self.ensure(q.shape[1])
        if q.shape[0] > sh.max_batch:
            raise _cabi.QreError(f"batch of {q.shape[0]} exceeds max_batch={sh.max_batch}")
        cfg.scan_mode = self.scan_mode
        if route_in is None:
            return sh.lookup_route(q, q_ctx, cfg, now_us)
        return sh.lookup_route(q, q_ctx, cfg, now_us, tok=route_in["tok"], heur_dev=route_in["heur_dev"],
                               heur_conf=route_in["heur_conf"])


    @_locked
    def scan_overlapped(self, q, q_ctx, now_us: int, cfg: _cabi.RouteConfig, route_in: Optional[dict],
                        host_work) -> dict:
        """`scan` with the caller's `host_work()` run on this thread while the GPU is busy with the
        batch (begin -> host_work -> end, all under the handle's lock): host bookkeeping that does
        not need the results — copying the embedding block the cache will keep — costs no wall time."""
        q = np.ascontiguousarray(q, np.float32)
        if q.ndim == 1:
            q = q.reshape(1, -1)
        sh = self.ensure(q.shape[1])
        if q.shape[0] > sh.max_batch:
            raise _cabi.QreError(f"batch of {q.shape[0]} exceeds max_batch={sh.max_batch}")
        cfg.scan_mode = self.scan_mode
        kw = {} if route_in is None else dict(tok=route_in["tok"], heur_dev=route_in["heur_dev"],
                                              heur_conf=route_in["heur_conf"])
        sh.lookup_route_begin(0, q, q_ctx, cfg, now_us, **kw)
        try:
            host_work()
        finally:
            out = sh.lookup_route_end(0)
        return out


class ShardedEngine:
    """The same store interface over SEVERAL GPUs driven by one process (`config["gpus"] = [0, 1, ..]`):
    one `Engine` per GPU, rows dealt round-robin, slots numbered `local_slot * n + shard` so the
    host classes keep addressing rows by one integer.  A scan is one `qre_group_lookup_route` call:
    every GPU scans and exactly re-scores its rows, the candidate lists meet on the first GPU and
    the merge + routing kernel runs there.  Decisions equal the one-GPU engine's (tested)."""

    def __init__(self, capacity: int = 512, devices: Sequence[int] = (0,), max_batch: int = 64,
                 dim: Optional[int] = None, scan_mode: int = _cabi.SCAN_TENSOR):
        self.devices = [int(d) for d in devices]
        if not self.devices or len(self.devices) > 16:
            raise ValueError("config['gpus'] must list 1..16 GPU indices")
        self.n = len(self.devices)
        self.capacity = max(int(capacity), 1)
        self.max_batch, self.scan_mode = int(max_batch), int(scan_mode)
        self._lock = threading.RLock()
        self._next = 0
        self._centroids = None
        self._parts = [Engine(self._per(), d, max_batch, None, scan_mode) for d in self.devices]
        if dim is not None:
            self.ensure(dim)

    def _per(self) -> int:
        return (self.capacity + self.n - 1) // self.n + 1

    # ---- lifetime ----
    @_locked
    def ensure(self, dim: int):
        for p in self._parts:
            p.capacity = max(p.capacity, self._per())
            p.ensure(dim)
        return self._parts[0].shard

    @property
    def shard(self):
        return self._parts[0].shard

    @_locked
    def close(self) -> None:
        for p in self._parts:
            p.close()

    @_locked
    def set_centroids(self, centroids, owner) -> None:
        self._centroids = np.ascontiguousarray(centroids, np.float32)
        for p in self._parts:
            p.set_centroids(centroids, owner)

    @_locked
    def set_perf_health(self, health) -> None:
        for p in self._parts:
            p.set_perf_health(health)

    @property
    def n_centroids(self) -> int:
        return self._parts[0].n_centroids

    # ---- rows: global slot = local * n + shard ----
    def _split(self, slots):
        slots = np.asarray(list(slots) if not isinstance(slots, np.ndarray) else slots, np.int64)
        return slots, slots % self.n, slots // self.n

    @_locked
    def append(self, row, ctx: int, ts_us: int, lru: int, pdev: int, pconf: float) -> int:
        d = self._next % self.n
        self._next += 1
        return self._parts[d].append(row, ctx, ts_us, lru, pdev, pconf) * self.n + d

    @_locked
    def append_many(self, rows, ctx, ts_us, lru, pdev, pconf) -> np.ndarray:
        rows = np.ascontiguousarray(rows, np.float32)
        k = rows.shape[0]
        dest = (self._next + np.arange(k)) % self.n
        self._next += k
        out = np.empty(k, np.int64)
        col = lambda v, sel: np.asarray(v)[sel]   # noqa: E731
        for d in range(self.n):
            sel = np.nonzero(dest == d)[0]
            if sel.size:
                loc = self._parts[d].append_many(rows[sel], col(ctx, sel), col(ts_us, sel), col(lru, sel),
                                                 col(pdev, sel), col(pconf, sel))
                out[sel] = np.asarray(loc, np.int64) * self.n + d
        return out

    @_locked
    def update(self, slot: int, row=None, ts_us=None, lru=None, pdev=None, pconf=None) -> None:
        self._parts[int(slot) % self.n].update(int(slot) // self.n, row, ts_us, lru, pdev, pconf)

    @_locked
    def update_many(self, slots, rows=None, ts_us=None, lru=None, pdev=None, pconf=None) -> None:
        slots, sh, loc = self._split(slots)
        pick = lambda v, sel: None if v is None else np.asarray(v)[sel]   # noqa: E731
        for d in range(self.n):
            sel = np.nonzero(sh == d)[0]
            if sel.size:
                self._parts[d].update_many(loc[sel], pick(rows, sel), pick(ts_us, sel), pick(lru, sel),
                                           pick(pdev, sel), pick(pconf, sel))

    @_locked
    def tombstone(self, slots) -> None:
        slots, sh, loc = self._split(slots)
        for d in range(self.n):
            sel = np.nonzero(sh == d)[0]
            if sel.size:
                self._parts[d].tombstone(loc[sel])

    @_locked
    def clear(self) -> None:
        for p in self._parts:
            p.clear()

    @_locked
    def read_rows(self, slots) -> np.ndarray:
        slots, sh, loc = self._split(slots)
        out = np.empty((slots.size, self._parts[0].shard.dim), np.float32)
        for d in range(self.n):
            sel = np.nonzero(sh == d)[0]
            if sel.size:
                out[sel] = self._parts[d].read_rows(loc[sel])
        return out

    @_locked
    def scan(self, q, q_ctx, now_us: int, cfg: _cabi.RouteConfig, route_in: Optional[dict] = None) -> dict:
        q = np.ascontiguousarray(q, np.float32)
        if q.ndim == 1:
            q = q.reshape(1, -1)
        self.ensure(q.shape[1])
        if q.shape[0] > self.max_batch:
            raise _cabi.QreError(f"batch of {q.shape[0]} exceeds max_batch={self.max_batch}")
        cfg.scan_mode = self.scan_mode
        shards = [p.shard for p in self._parts]
        ri = route_in or {}
        out = _cabi.group_lookup_route(shards, q, q_ctx, cfg, now_us, tok=ri.get("tok"),
                                       heur_dev=ri.get("heur_dev"), heur_conf=ri.get("heur_conf"))
        hit = out["hit_slot"] >= 0
        out["hit_slot"] = np.where(hit, out["hit_slot"] * self.n + out["hit_shard"].astype(np.int64), -1)
        return out


