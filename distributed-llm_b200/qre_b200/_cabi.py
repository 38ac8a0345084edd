"""ctypes binding of libqre_b200.so (include/qre_b200.h).

This is the only place the Python host code touches native code.  There is no CPU fallback:
if the library cannot be loaded, or no sm_100 GPU is present, constructing a `Shard` raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

TOPK = 8
MAX_DEVICES = 32
MAX_CENTROIDS = 256
CTX_TOMBSTONE = 0xFFFFFFFFFFFFFFFF
SCAN_TENSOR, SCAN_EXACT = 0, 1
FLAG_ZERO_NORM, FLAG_OVERFLOW = 1, 2
SEM_METHOD_NAMES = ("semantic", "semantic_fallback_irrelevant", "semantic_fallback_ambiguous")

CAND_DTYPE = np.dtype([("sim", "<f4"), ("shard", "<u4"), ("slot", "<i8"), ("lru_seq", "<u8"),
                       ("pred_conf", "<f8")])
assert CAND_DTYPE.itemsize == 32

EXPORTS = (
    "qre_abi_version", "qre_last_error", "qre_device_count", "qre_cache_create", "qre_cache_destroy",
    "qre_cache_info", "qre_cache_stream", "qre_cache_append", "qre_cache_append_dev", "qre_cache_update",
    "qre_cache_tombstone", "qre_cache_clear", "qre_cache_read_rows", "qre_set_centroids",
    "qre_set_perf_health", "qre_lookup_route", "qre_scan_dev", "qre_resolve_overflow_dev",
    "qre_merge_route_dev", "qre_launch_count", "qre_debug_gemm", "qre_profile", "qre_profile_read",
    "qre_exact_lookup_noalloc", "qre_cache_snapshot_save", "qre_cache_snapshot_load",
    "qre_lookup_route_begin", "qre_lookup_route_end", "qre_host_features", "qre_host_features_packed",
    "qre_submit_dev", "qre_wait_dev", "qre_group_lookup_route", "qre_xchg_init", "qre_xchg_connect", "qre_scan_publish_dev", "qre_merge_route_xchg_dev", "qre_publish_dev",
    # include/qre_encoder_b200.h
    "qre_encoder_create", "qre_encoder_destroy", "qre_encoder_set_embeddings", "qre_encoder_set_layer",
    "qre_encode", "qre_encode_dev", "qre_encoder_launch_count", "qre_encoder_debug_gemm",
    "qre_wordpiece_create", "qre_wordpiece_destroy", "qre_wordpiece_encode", "qre_wordpiece_encode_packed",
)


class QreError(RuntimeError):
    pass


class Params(C.Structure):
    _fields_ = [
        ("similarity_threshold", C.c_double), ("now_us", C.c_int64), ("ttl_us", C.c_int64),
        ("scan_mode", C.c_int32), ("do_route", C.c_int32), ("n_devices", C.c_int32),
        ("token_threshold", C.c_int32), ("w_token", C.c_double), ("w_semantic", C.c_double),
        ("w_heuristic", C.c_double), ("w_perf", C.c_double), ("sem_min_similarity", C.c_double),
        ("sem_margin_threshold", C.c_double), ("pipeline_slot", C.c_int32), ("reserve_sms", C.c_int32),
        ("slice_results", C.c_int32), ("reserved", C.c_int32),
    ]


RESULT_FIELDS = ("hit_slot", "hit_sim", "hit_pred_dev", "hit_pred_conf", "hit_shard", "cand",
                 "centroid_sim", "tok_dev", "tok_conf", "sem_dev", "sem_conf", "sem_method", "dev",
                 "conf", "score", "flags")


class Result(C.Structure):
    _fields_ = [(f, C.c_void_p) for f in RESULT_FIELDS]


def lib_path() -> str:
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib",
                        "libqre_b200.so")


_LIB = None


def load() -> C.CDLL:
    """Load the shared library (building nothing: see qre_b200.build / __graft_entry__.build)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = os.environ.get("QRE_B200_LIB", lib_path())
    if not os.path.isfile(path):
        raise QreError(f"{path} not found: build it with `python distributed-llm_b200/qre_b200/build.py` "
                       "(there is no CPU fallback)")
    lib = C.CDLL(path)
    lib.qre_last_error.restype = C.c_char_p
    for name in EXPORTS:
        getattr(lib, name)
    _LIB = lib
    return lib


def _check(lib, rc: int) -> None:
    if rc != 0:
        raise QreError(f"libqre_b200 error {rc}: {lib.qre_last_error().decode()}")


def _ptr(a: Optional[np.ndarray]):
    # ndarray.ctypes builds a helper object per call (~1.5 us); the array interface is cheaper and
    # this sits on the latency path of every small lookup (21 pointers per call)
    return None if a is None else C.c_void_p(a.__array_interface__["data"][0])


def _arr(x, dtype, shape=None) -> np.ndarray:
    a = np.ascontiguousarray(x, dtype=dtype)
    if shape is not None and tuple(a.shape) != tuple(shape):
        raise ValueError(f"expected shape {shape}, got {a.shape}")
    return a


@dataclass
class RouteConfig:
    """Knobs of one lookup+route pass; defaults = the reference's shipped configs
    (query_router_engine.py:704-731)."""
    similarity_threshold: float = 0.85
    ttl_us: int = 3600 * 1_000_000
    n_devices: int = 2
    token_threshold: int = 1000
    w_token: float = 0.25
    w_semantic: float = 0.45
    w_heuristic: float = 0.30
    w_perf: float = 0.0
    sem_min_similarity: float = 0.05
    sem_margin_threshold: float = 0.03
    scan_mode: int = SCAN_TENSOR
    pipeline_slot: int = 0
    reserve_sms: int = 0
    slice_results: int = 0

    def to_c(self, now_us: int, do_route: bool) -> Params:
        return Params(self.similarity_threshold, int(now_us), int(self.ttl_us), self.scan_mode,
                      1 if do_route else 0, self.n_devices, self.token_threshold, self.w_token,
                      self.w_semantic, self.w_heuristic, self.w_perf, self.sem_min_similarity,
                      self.sem_margin_threshold, self.pipeline_slot, self.reserve_sms, self.slice_results, 0)


class Shard:
    """One GPU's slice of the routing cache (a `qre_cache_t` handle)."""

    def __init__(self, dim: int, capacity: int, max_batch: int = 4096, device: int = 0):
        self.lib = load()
        self.dim, self.capacity, self.max_batch, self.device = int(dim), int(capacity), int(max_batch), int(device)
        h = C.c_void_p()
        _check(self.lib, self.lib.qre_cache_create(C.c_int(device), C.c_int(dim), C.c_int64(capacity),
                                                   C.c_int(max_batch), C.byref(h)))
        self.h = h
        self.n_centroids = 0

    def close(self) -> None:
        if getattr(self, "h", None) is not None and self.h:
            self.lib.qre_cache_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- info ----
    def info(self) -> dict:
        cap, hwm, live = C.c_int64(), C.c_int64(), C.c_int64()
        dim, mb, dev = C.c_int(), C.c_int(), C.c_int()
        _check(self.lib, self.lib.qre_cache_info(self.h, C.byref(cap), C.byref(hwm), C.byref(live),
                                                 C.byref(dim), C.byref(mb), C.byref(dev)))
        return dict(capacity=cap.value, high_water=hwm.value, live=live.value, dim=dim.value,
                    max_batch=mb.value, device=dev.value)

    @property
    def stream(self) -> int:
        s = C.c_void_p()
        _check(self.lib, self.lib.qre_cache_stream(self.h, C.byref(s)))
        return s.value or 0

    def launch_count(self) -> int:
        n = C.c_int64()
        _check(self.lib, self.lib.qre_launch_count(self.h, C.byref(n)))
        return n.value

    # ---- rows ----
    def append(self, rows, ctx, ts_us, lru_seq, pred_dev, pred_conf) -> np.ndarray:
        rows = _arr(rows, np.float32)
        n = rows.shape[0]
        if rows.ndim != 2 or rows.shape[1] != self.dim:
            raise ValueError(f"rows must be [n, {self.dim}]")
        ctx, ts_us = _arr(ctx, np.uint64, (n,)), _arr(ts_us, np.int64, (n,))
        lru_seq, pred_dev = _arr(lru_seq, np.uint64, (n,)), _arr(pred_dev, np.uint8, (n,))
        pred_conf = _arr(pred_conf, np.float64, (n,))
        slots = np.empty(n, np.int64)
        _check(self.lib, self.lib.qre_cache_append(self.h, C.c_int64(n), _ptr(rows), _ptr(ctx), _ptr(ts_us),
                                                   _ptr(lru_seq), _ptr(pred_dev), _ptr(pred_conf), _ptr(slots)))
        return slots

    def append_dev(self, n, rows_ptr, ctx_ptr, ts_ptr, lru_ptr, pdev_ptr, pconf_ptr, stream=0) -> np.ndarray:
        slots = np.empty(n, np.int64)
        _check(self.lib, self.lib.qre_cache_append_dev(
            self.h, C.c_int64(n), C.c_void_p(rows_ptr), C.c_void_p(ctx_ptr), C.c_void_p(ts_ptr),
            C.c_void_p(lru_ptr), C.c_void_p(pdev_ptr), C.c_void_p(pconf_ptr), _ptr(slots), C.c_void_p(stream)))
        return slots

    def update(self, slots, rows=None, ctx=None, ts_us=None, lru_seq=None, pred_dev=None, pred_conf=None):
        slots = _arr(slots, np.int64)
        n = slots.shape[0]
        rows = None if rows is None else _arr(rows, np.float32, (n, self.dim))
        ctx = None if ctx is None else _arr(ctx, np.uint64, (n,))
        ts_us = None if ts_us is None else _arr(ts_us, np.int64, (n,))
        lru_seq = None if lru_seq is None else _arr(lru_seq, np.uint64, (n,))
        pred_dev = None if pred_dev is None else _arr(pred_dev, np.uint8, (n,))
        pred_conf = None if pred_conf is None else _arr(pred_conf, np.float64, (n,))
        _check(self.lib, self.lib.qre_cache_update(self.h, C.c_int64(n), _ptr(slots), _ptr(rows), _ptr(ctx),
                                                   _ptr(ts_us), _ptr(lru_seq), _ptr(pred_dev), _ptr(pred_conf)))

    def tombstone(self, slots) -> None:
        slots = _arr(slots, np.int64)
        _check(self.lib, self.lib.qre_cache_tombstone(self.h, C.c_int64(slots.shape[0]), _ptr(slots)))

    def clear(self) -> None:
        _check(self.lib, self.lib.qre_cache_clear(self.h))

    def read_rows(self, slots) -> np.ndarray:
        slots = _arr(slots, np.int64)
        out = np.empty((slots.shape[0], self.dim), np.float32)
        _check(self.lib, self.lib.qre_cache_read_rows(self.h, C.c_int64(slots.shape[0]), _ptr(slots), _ptr(out)))
        return out

    def snapshot_save(self, path: str) -> None:
        _check(self.lib, self.lib.qre_cache_snapshot_save(self.h, C.c_char_p(os.fsencode(path))))

    def snapshot_load(self, path: str) -> None:
        _check(self.lib, self.lib.qre_cache_snapshot_load(self.h, C.c_char_p(os.fsencode(path))))

    # ---- router state ----
    def set_centroids(self, centroids, owner) -> None:
        c = _arr(centroids, np.float32)
        if c.ndim != 2 or c.shape[1] != self.dim:
            raise ValueError(f"centroids must be [K, {self.dim}]")
        o = _arr(owner, np.uint8, (c.shape[0],))
        _check(self.lib, self.lib.qre_set_centroids(self.h, C.c_int(c.shape[0]), _ptr(c), _ptr(o)))
        self.n_centroids = c.shape[0]

    def set_perf_health(self, health) -> None:
        hlt = _arr(health, np.float64)
        _check(self.lib, self.lib.qre_set_perf_health(self.h, C.c_int(hlt.shape[0]), _ptr(hlt)))

    # ---- the hot path, host buffers ----
    def _host_outputs(self, m: int, do_route: bool, D: int, want_cand: bool) -> dict:
        K = self.n_centroids
        out = dict(hit_slot=np.empty(m, np.int64), hit_sim=np.empty(m, np.float32),
                   hit_pred_dev=np.empty(m, np.uint8), hit_pred_conf=np.empty(m, np.float64),
                   hit_shard=np.empty(m, np.uint32), flags=np.empty(m, np.uint8))
        if want_cand:
            out["cand"] = np.empty((m, TOPK), CAND_DTYPE)
        if do_route:
            out.update(centroid_sim=np.empty((m, K), np.float32), tok_dev=np.empty(m, np.uint8),
                       tok_conf=np.empty(m, np.float64), sem_dev=np.empty(m, np.uint8),
                       sem_conf=np.empty(m, np.float64), sem_method=np.empty(m, np.uint8),
                       dev=np.empty(m, np.uint8), conf=np.empty(m, np.float64),
                       score=np.empty((m, D), np.float64))
        return out

    def lookup_route_begin(self, slot: int, queries, q_ctx, cfg: RouteConfig, now_us: int, tok=None,
                           heur_dev=None, heur_conf=None, want_cand: bool = False) -> None:
        """Enqueue one batch on host slot 0 / 1 / 2 and return; pair with `lookup_route_end(slot)`.
        The arrays must stay alive and unchanged until then (they are kept referenced here)."""
        q = _arr(queries, np.float32)
        m = q.shape[0]
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise ValueError(f"queries must be [m, {self.dim}]")
        q_ctx = _arr(q_ctx, np.uint64, (m,))
        do_route = tok is not None
        if do_route:
            tok = _arr(tok, np.int32, (m,))
            heur_dev = _arr(heur_dev, np.uint8, (m,))
            heur_conf = _arr(heur_conf, np.float64, (m,))
        p = cfg.to_c(now_us, do_route)
        _check(self.lib, self.lib.qre_lookup_route_begin(self.h, C.c_int(slot), C.c_int(m), _ptr(q), _ptr(q_ctx),
                                                         _ptr(tok), _ptr(heur_dev), _ptr(heur_conf), C.byref(p)))
        if not hasattr(self, "_inflight"):
            self._inflight = {}
        self._inflight[slot] = (m, do_route, cfg.n_devices, want_cand, (q, q_ctx, tok, heur_dev, heur_conf))

    def lookup_route_end(self, slot: int, out: Optional[dict] = None) -> dict:
        """Wait for the batch begun on `slot` and return its result arrays.  `out`: a dict returned
        by an earlier `lookup_route_end` for a batch of the same shape, to be filled again instead of
        allocating fresh arrays (the C call writes into the caller's arrays either way)."""
        m, do_route, D, want_cand, _keep = self._inflight.pop(slot)
        if out is None or out["hit_slot"].shape[0] != m or ("dev" in out) != do_route or ("cand" in out) != want_cand \
                or (do_route and out["score"].shape[1] != D):
            out = self._host_outputs(m, do_route, D, want_cand)
        res = Result(**{f: _ptr(out.get(f)) for f in RESULT_FIELDS})
        _check(self.lib, self.lib.qre_lookup_route_end(self.h, C.c_int(slot), C.byref(res)))
        return out

    def lookup_route(self, queries, q_ctx, cfg: RouteConfig, now_us: int, tok=None, heur_dev=None,
                     heur_conf=None, want_cand: bool = False) -> dict:
        q = _arr(queries, np.float32)
        m = q.shape[0]
        if q.ndim != 2 or q.shape[1] != self.dim:
            raise ValueError(f"queries must be [m, {self.dim}]")
        q_ctx = _arr(q_ctx, np.uint64, (m,))
        do_route = tok is not None
        if do_route:
            tok = _arr(tok, np.int32, (m,))
            heur_dev = _arr(heur_dev, np.uint8, (m,))
            heur_conf = _arr(heur_conf, np.float64, (m,))
        out = self._host_outputs(m, do_route, cfg.n_devices, want_cand)
        res = Result(**{f: _ptr(out.get(f)) for f in RESULT_FIELDS})
        p = cfg.to_c(now_us, do_route)
        _check(self.lib, self.lib.qre_lookup_route(self.h, C.c_int(m), _ptr(q), _ptr(q_ctx), _ptr(tok),
                                                   _ptr(heur_dev), _ptr(heur_conf), C.byref(p), C.byref(res)))
        return out

    # ---- the hot path, device pointers (integers from torch .data_ptr()) ----
    def scan_dev(self, m, q_ptr, q_ctx_ptr, cfg: RouteConfig, now_us, shard, cand_ptr, ovf_ptr=0, stream=0):
        p = cfg.to_c(now_us, False)
        _check(self.lib, self.lib.qre_scan_dev(self.h, C.c_int(m), C.c_void_p(q_ptr), C.c_void_p(q_ctx_ptr),
                                               C.byref(p), C.c_int(shard), C.c_void_p(cand_ptr),
                                               C.c_void_p(ovf_ptr), C.c_void_p(stream)))

    def resolve_overflow_dev(self, m, cfg: RouteConfig, now_us, shard, cand_ptr, stream=0):
        p = cfg.to_c(now_us, False)
        _check(self.lib, self.lib.qre_resolve_overflow_dev(self.h, C.c_int(m), C.byref(p), C.c_int(shard),
                                                           C.c_void_p(cand_ptr), C.c_void_p(stream)))

    def merge_route_dev(self, m, n_shards, cand_all_ptr, cfg: RouteConfig, now_us, out_ptrs: dict,
                        tok_ptr=0, heur_dev_ptr=0, heur_conf_ptr=0, stream=0):
        do_route = bool(tok_ptr)
        p = cfg.to_c(now_us, do_route)
        res = Result(**{f: C.c_void_p(out_ptrs.get(f, 0) or None) for f in RESULT_FIELDS})
        _check(self.lib, self.lib.qre_merge_route_dev(
            self.h, C.c_int(m), C.c_int(n_shards), C.c_void_p(cand_all_ptr), C.c_void_p(tok_ptr or None),
            C.c_void_p(heur_dev_ptr or None), C.c_void_p(heur_conf_ptr or None), C.byref(p), C.byref(res),
            C.c_void_p(stream)))

    def submit_dev(self, slot, m, q_ptr, q_ctx_ptr, cfg: RouteConfig, now_us, out_ptrs: dict, tok_ptr=0,
                   heur_dev_ptr=0, heur_conf_ptr=0, in_stream=0):
        """qre_submit_dev: one batch through the library's own three-stream pipeline (slot 0/1)."""
        p = cfg.to_c(now_us, bool(tok_ptr))
        res = Result(**{f: C.c_void_p(out_ptrs.get(f, 0) or None) for f in RESULT_FIELDS})
        _check(self.lib, self.lib.qre_submit_dev(
            self.h, C.c_int(slot), C.c_int(m), C.c_void_p(q_ptr), C.c_void_p(q_ctx_ptr), C.c_void_p(tok_ptr or None),
            C.c_void_p(heur_dev_ptr or None), C.c_void_p(heur_conf_ptr or None), C.byref(p), C.byref(res),
            C.c_void_p(in_stream or None)))

    def submit_dev_prepared(self, slot, m, q_ptr, q_ctx_ptr, cfg: RouteConfig, out_ptrs: dict, tok_ptr=0,
                            heur_dev_ptr=0, heur_conf_ptr=0):
        """The same call with every ctypes object built ONCE: returns `fn(now_us, in_stream)`.  At 8 GPUs a
        batch lasts ~250 us, so the ~100 us of per-call argument marshalling of `submit_dev` (15 pointer
        look-ups, two structs) would leave the GPU waiting for the host."""
        p = cfg.to_c(0, bool(tok_ptr))
        res = Result(**{f: C.c_void_p(out_ptrs.get(f, 0) or None) for f in RESULT_FIELDS})
        args = (self.h, C.c_int(slot), C.c_int(m), C.c_void_p(q_ptr), C.c_void_p(q_ctx_ptr),
                C.c_void_p(tok_ptr or None), C.c_void_p(heur_dev_ptr or None), C.c_void_p(heur_conf_ptr or None),
                C.byref(p), C.byref(res))
        fn, lib = self.lib.qre_submit_dev, self.lib

        def call(now_us: int, in_stream: int = 0, _keep=(p, res)):
            p.now_us = now_us
            rc = fn(*args, C.c_void_p(in_stream or None))
            if rc != 0:
                _check(lib, rc)
        return call

    def wait_dev(self, slot, stream=0):
        _check(self.lib, self.lib.qre_wait_dev(self.h, C.c_int(slot), C.c_void_p(stream or None)))

    def profile(self, enable: bool) -> None:
        _check(self.lib, self.lib.qre_profile(self.h, C.c_int(1 if enable else 0)))

    def profile_read(self):
        ms, n = C.c_double(), C.c_int()
        _check(self.lib, self.lib.qre_profile_read(self.h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- peer exchange ----
    def xchg_init(self, world: int, rank: int) -> bytes:
        buf = C.create_string_buffer(64)
        _check(self.lib, self.lib.qre_xchg_init(self.h, C.c_int(world), C.c_int(rank), buf))
        return buf.raw

    def xchg_connect(self, peer: int, handle: bytes) -> None:
        _check(self.lib, self.lib.qre_xchg_connect(self.h, C.c_int(peer), C.c_char_p(handle)))

    def publish_dev(self, m, cand_ptr, stream=0):
        _check(self.lib, self.lib.qre_publish_dev(self.h, C.c_int(m), C.c_void_p(cand_ptr), C.c_void_p(stream)))

    def scan_publish_dev(self, m, q_ptr, q_ctx_ptr, cfg: RouteConfig, now_us, ovf_ptr=0, stream=0):
        p = cfg.to_c(now_us, False)
        _check(self.lib, self.lib.qre_scan_publish_dev(self.h, C.c_int(m), C.c_void_p(q_ptr), C.c_void_p(q_ctx_ptr),
                                                       C.byref(p), C.c_void_p(ovf_ptr or None), C.c_void_p(stream)))

    def merge_route_xchg_dev(self, m, cfg: RouteConfig, now_us, out_ptrs: dict, tok_ptr=0, heur_dev_ptr=0,
                             heur_conf_ptr=0, stream=0):
        p = cfg.to_c(now_us, bool(tok_ptr))
        res = Result(**{f: C.c_void_p(out_ptrs.get(f, 0) or None) for f in RESULT_FIELDS})
        _check(self.lib, self.lib.qre_merge_route_xchg_dev(
            self.h, C.c_int(m), C.c_void_p(tok_ptr or None), C.c_void_p(heur_dev_ptr or None),
            C.c_void_p(heur_conf_ptr or None), C.byref(p), C.byref(res), C.c_void_p(stream)))

    def debug_gemm(self, m: int, row0: int, n: int) -> np.ndarray:
        out = np.empty((m, n), np.float32)
        _check(self.lib, self.lib.qre_debug_gemm(self.h, C.c_int(m), C.c_int64(row0), C.c_int(n), _ptr(out)))
        return out


def group_lookup_route(shards, queries, q_ctx, cfg: RouteConfig, now_us: int, tok=None, heur_dev=None,
                       heur_conf=None, want_cand: bool = False) -> dict:
    """qre_group_lookup_route over a list of `Shard`s (one per GPU, one process): the row-sharded
    lookup + route with the merge on the first shard's GPU.  Same outputs as `Shard.lookup_route`;
    `hit_shard` indexes `shards`."""
    first = shards[0]
    q = _arr(queries, np.float32)
    m = q.shape[0]
    if q.ndim != 2 or q.shape[1] != first.dim:
        raise ValueError(f"queries must be [m, {first.dim}]")
    q_ctx = _arr(q_ctx, np.uint64, (m,))
    do_route = tok is not None
    if do_route:
        tok, heur_dev = _arr(tok, np.int32, (m,)), _arr(heur_dev, np.uint8, (m,))
        heur_conf = _arr(heur_conf, np.float64, (m,))
    out = first._host_outputs(m, do_route, cfg.n_devices, want_cand)
    res = Result(**{f: _ptr(out.get(f)) for f in RESULT_FIELDS})
    p = cfg.to_c(now_us, do_route)
    hs = (C.c_void_p * len(shards))(*[s.h for s in shards])
    _check(first.lib, first.lib.qre_group_lookup_route(hs, C.c_int(len(shards)), C.c_int(m), _ptr(q), _ptr(q_ctx),
                                                       _ptr(tok), _ptr(heur_dev), _ptr(heur_conf), C.byref(p),
                                                       C.byref(res)))
    return out


NOALLOC_MAX_QUERIES, NOALLOC_MAX_DIM = 8, 1024


def exact_lookup_noalloc(device: int, n_rows: int, dim: int, rows_ptr: int, row_ctx_ptr: int, row_ts_ptr: int,
                         row_lru_ptr: int, m: int, q_ptr: int, q_ctx_ptr: int, threshold: float, now_us: int,
                         ttl_us: int, hit_row_ptr: int, hit_sim_ptr: int, stream: int = 0) -> None:
    """Thin call of qre_exact_lookup_noalloc (include/qre_b200.h): every array is a DEVICE pointer
    the caller owns; nothing is allocated on the way.  Raises QreError with the library's message
    when a size limit is exceeded or a CUDA call fails."""
    lib = load()
    err = C.create_string_buffer(256)
    rc = lib.qre_exact_lookup_noalloc(
        C.c_int(device), C.c_int64(n_rows), C.c_int(dim), C.c_void_p(rows_ptr), C.c_void_p(row_ctx_ptr),
        C.c_void_p(row_ts_ptr), C.c_void_p(row_lru_ptr), C.c_int(m), C.c_void_p(q_ptr), C.c_void_p(q_ctx_ptr),
        C.c_double(threshold), C.c_int64(now_us), C.c_int64(ttl_us), C.c_void_p(hit_row_ptr),
        C.c_void_p(hit_sim_ptr), C.c_void_p(stream or None), err, C.c_size_t(len(err)))
    if rc != 0:
        raise QreError(f"libqre_b200 error {rc}: {err.value.decode()}")


class HeurParams(C.Structure):
    _fields_ = [("long_chars", C.c_int32), ("multi_qmarks", C.c_int32), ("code_markers_needed", C.c_int32),
                ("context_chars", C.c_int32), ("token_threshold", C.c_int32), ("n_devices", C.c_int32)]


def host_features(queries, context_lens, hp: HeurParams, n_threads: int = 0) -> dict:
    """qre_host_features for a list of str: token estimate + heuristic rule per query, computed on
    C++ threads (no GIL).  rule == -1 marks a query the caller must classify itself
    (malformed UTF-8: cannot happen for a Python str)."""
    lib = load()
    n = len(queries)
    enc = [q.encode("utf-8") for q in queries]
    ptrs = (C.c_char_p * n)(*enc)
    qlen = np.array([len(b) for b in enc], np.int64)
    clen = np.ascontiguousarray(context_lens, np.int64)
    out = dict(tok=np.empty(n, np.int32), heur_dev=np.empty(n, np.uint8), heur_conf=np.empty(n, np.float64),
               rule=np.empty(n, np.int32), cat=np.empty(n, np.int32))
    if n_threads <= 0:
        n_threads = min(32, os.cpu_count() or 1)
    rc = lib.qre_host_features(C.c_int64(n), ptrs, _ptr(qlen), _ptr(clen), C.byref(hp), C.c_int(n_threads),
                               _ptr(out["tok"]), _ptr(out["heur_dev"]), _ptr(out["heur_conf"]), _ptr(out["rule"]),
                               _ptr(out["cat"]))
    if rc != 0:
        raise QreError(f"qre_host_features failed with {rc}")
    return out


def host_features_packed(queries, context_lens, hp: HeurParams, context_keys=None, n_threads: int = 0) -> dict:
    """`host_features` without per-query marshalling: the queries travel as one UTF-8 blob plus
    offsets (qre_host_features_packed).  With `context_keys` (one per query) the result also
    carries `keys`: the cache's exact-match md5 key of every query (None where rule == -1, i.e. a
    query handed back, which the caller hashes itself)."""
    lib = load()
    n = len(queries)
    if n == 0:
        out = dict(tok=np.empty(0, np.int32), heur_dev=np.empty(0, np.uint8), heur_conf=np.empty(0, np.float64),
                   rule=np.empty(0, np.int32), cat=np.empty(0, np.int32))
        if context_keys is not None:
            out["keys"] = []
        return out
    blob = "".join(queries).encode("utf-8")
    lens = np.fromiter(map(len, queries), np.int64, n)
    if int(lens.sum()) != len(blob):                 # some query is not ASCII: byte lengths differ
        lens = np.fromiter((len(q.encode("utf-8")) for q in queries), np.int64, n)
    off = np.zeros(n + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    clen = np.ascontiguousarray(context_lens, np.int64)
    out = dict(tok=np.empty(n, np.int32), heur_dev=np.empty(n, np.uint8), heur_conf=np.empty(n, np.float64),
               rule=np.empty(n, np.int32), cat=np.empty(n, np.int32))
    if n_threads <= 0:
        n_threads = min(32, os.cpu_count() or 1)
    cblob, coff, cidx, n_ctx, hexbuf = None, None, None, 0, None
    if context_keys is not None:
        table: dict = {}
        cidx = np.fromiter((table.setdefault(k, len(table)) for k in context_keys), np.int32, n)
        enc = [k.encode("utf-8") for k in table]
        cblob = b"".join(enc)
        coff = np.zeros(len(enc) + 1, np.int64)
        np.cumsum([len(b) for b in enc], out=coff[1:])
        n_ctx = len(enc)
        hexbuf = np.zeros(32 * n, np.uint8)
    rc = lib.qre_host_features_packed(
        C.c_int64(n), C.c_char_p(blob), _ptr(off), _ptr(clen), C.byref(hp), C.c_int(n_threads),
        _ptr(out["tok"]), _ptr(out["heur_dev"]), _ptr(out["heur_conf"]), _ptr(out["rule"]), _ptr(out["cat"]),
        C.c_char_p(cblob) if cblob is not None else None, _ptr(coff) if coff is not None else None,
        _ptr(cidx) if cidx is not None else None, C.c_int32(n_ctx), _ptr(hexbuf) if hexbuf is not None else None)
    if rc != 0:
        raise QreError(f"qre_host_features_packed failed with {rc}")
    if hexbuf is not None:
        hx = hexbuf.tobytes().decode("latin-1")
        keys = [hx[i:i + 32] for i in range(0, 32 * n, 32)]
        for i in np.nonzero(out["rule"] < 0)[0].tolist():
            keys[i] = None
        out["keys"] = keys
    return out


# ---- include/qre_encoder_b200.h -----------------------------------------------------------------
class EncoderConfig(C.Structure):
    """qre_encoder_config_t; the defaults are all-MiniLM-L6-v2 (qre:125, 510)."""
    _fields_ = [("vocab_size", C.c_int32), ("hidden", C.c_int32), ("layers", C.c_int32), ("heads", C.c_int32),
                ("intermediate", C.c_int32), ("max_positions", C.c_int32), ("type_vocab", C.c_int32),
                ("ln_eps", C.c_float), ("max_seq_len", C.c_int32), ("reserved", C.c_int32)]

    def __init__(self, vocab_size=30522, hidden=384, layers=6, heads=12, intermediate=1536, max_positions=512,
                 type_vocab=2, ln_eps=1e-12, max_seq_len=256):
        super().__init__(vocab_size, hidden, layers, heads, intermediate, max_positions, type_vocab, ln_eps,
                         max_seq_len, 0)


_LAYER_FIELDS = ("wq", "bq", "wk", "bk", "wv", "bv", "wo", "bo", "ln1_g", "ln1_b", "w1", "b1", "w2", "b2",
                 "ln2_g", "ln2_b")
_LAYER_KEYS = ("attention.self.query.weight", "attention.self.query.bias", "attention.self.key.weight",
               "attention.self.key.bias", "attention.self.value.weight", "attention.self.value.bias",
               "attention.output.dense.weight", "attention.output.dense.bias", "attention.output.LayerNorm.weight",
               "attention.output.LayerNorm.bias", "intermediate.dense.weight", "intermediate.dense.bias",
               "output.dense.weight", "output.dense.bias", "output.LayerNorm.weight", "output.LayerNorm.bias")


class EncoderLayer(C.Structure):
    _fields_ = [(f, C.c_void_p) for f in _LAYER_FIELDS]


class Encoder:
    """One on-GPU query embedder (qre_encoder_t).  Weights: a dict of float32 arrays under
    transformers' `BertModel.state_dict()` names."""

    def __init__(self, cfg: EncoderConfig, weights: dict, max_tokens: int = 1 << 18, max_seqs: int = 4096,
                 device: int = 0):
        self._lib = load()
        self._h = C.c_void_p()
        self.cfg, self.device, self.max_tokens, self.max_seqs = cfg, device, int(max_tokens), int(max_seqs)
        _check(self._lib, self._lib.qre_encoder_create(C.c_int(device), C.byref(cfg), C.c_int64(max_tokens),
                                                       C.c_int(max_seqs), C.byref(self._h)))
        try:
            self.set_weights(weights)
        except Exception:
            self.close()
            raise

    def set_weights(self, w: dict) -> None:
        c, H, F = self.cfg, self.cfg.hidden, self.cfg.intermediate
        g = lambda k, shape: _arr(w[k], np.float32, shape)   # noqa: E731
        emb = [g("embeddings.word_embeddings.weight", (c.vocab_size, H)),
               g("embeddings.position_embeddings.weight", (c.max_positions, H)),
               g("embeddings.token_type_embeddings.weight", (c.type_vocab, H)),
               g("embeddings.LayerNorm.weight", (H,)), g("embeddings.LayerNorm.bias", (H,))]
        _check(self._lib, self._lib.qre_encoder_set_embeddings(self._h, *[_ptr(a) for a in emb]))
        shapes = ((H, H), (H,), (H, H), (H,), (H, H), (H,), (H, H), (H,), (H,), (H,), (F, H), (F,), (H, F), (H,),
                  (H,), (H,))
        for l in range(c.layers):
            arrs = [g(f"encoder.layer.{l}.{k}", sh) for k, sh in zip(_LAYER_KEYS, shapes)]
            lay = EncoderLayer(*[a.__array_interface__["data"][0] for a in arrs])
            _check(self._lib, self._lib.qre_encoder_set_layer(self._h, C.c_int(l), C.byref(lay)))

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.qre_encoder_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode_ids(self, seqs, normalize: bool = True) -> np.ndarray:
        """seqs: list of int id arrays -> float32 [m, hidden] (host in, host out)."""
        m = len(seqs)
        lens = np.fromiter((len(s) for s in seqs), np.int64, m)
        off = np.zeros(m + 1, np.int32)
        np.cumsum(lens, out=off[1:])
        ids = (np.concatenate([np.asarray(s, np.int32) for s in seqs]) if m and off[-1] else np.zeros(0, np.int32))
        return self.encode_packed(ids, off, normalize)

    def encode_packed(self, ids: np.ndarray, offsets: np.ndarray, normalize: bool = True) -> np.ndarray:
        ids, offsets = _arr(ids, np.int32), _arr(offsets, np.int32)
        m = offsets.shape[0] - 1
        out = np.empty((m, self.cfg.hidden), np.float32)
        _check(self._lib, self._lib.qre_encode(self._h, C.c_int(m), _ptr(ids) if ids.size else None, _ptr(offsets),
                                               C.c_int(int(normalize)), _ptr(out)))
        return out

    def encode_dev(self, m: int, ids_ptr: int, offsets_ptr: int, n_tokens: int, max_len: int, out_ptr: int,
                   normalize: bool = True, stream: int = 0) -> None:
        _check(self._lib, self._lib.qre_encode_dev(self._h, C.c_int(m), C.c_void_p(ids_ptr), C.c_void_p(offsets_ptr),
                                                   C.c_int64(n_tokens), C.c_int(max_len), C.c_int(int(normalize)),
                                                   C.c_void_p(out_ptr), C.c_void_p(stream)))

    def launch_count(self) -> int:
        n = C.c_int64()
        _check(self._lib, self._lib.qre_encoder_launch_count(self._h, C.byref(n)))
        return n.value


def encoder_debug_gemm(a, w, bias, residual=None, epilogue: int = 0, device: int = 0, ln_g=None, ln_b=None,
                       ln_eps: float = 0.0) -> np.ndarray:
    a, w, bias = _arr(a, np.float32), _arr(w, np.float32), _arr(bias, np.float32)
    rows, k = a.shape
    n = w.shape[0]
    res = None if residual is None else _arr(residual, np.float32, (rows, n))
    out = np.empty((rows, n), np.float32)
    lib = load()
    g = None if ln_g is None else _arr(ln_g, np.float32, (n,))
    b = None if ln_b is None else _arr(ln_b, np.float32, (n,))
    _check(lib, lib.qre_encoder_debug_gemm(C.c_int(device), C.c_int(rows), C.c_int(n), C.c_int(k), _ptr(a), _ptr(w),
                                           _ptr(bias), _ptr(res), C.c_int(epilogue), _ptr(g), _ptr(b),
                                           C.c_float(ln_eps), _ptr(out)))
    return out


class WordPiece:
    """Host-thread WordPiece (qre_wordpiece_t), ASCII and general Unicode texts; the rare text it hands back
    (a character whose tokenisation depends on its neighbours or differs between the reference tokenizers,
    an unassigned code point) has length -1."""

    def __init__(self, vocab_path: str, do_lower_case: bool = True):
        self._lib = load()
        self._h = C.c_void_p()
        _check(self._lib, self._lib.qre_wordpiece_create(vocab_path.encode(), C.c_int(int(do_lower_case)),
                                                         C.byref(self._h)))

    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.qre_wordpiece_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def encode(self, texts, max_len: int = 256, n_threads: int = 0):
        """-> (ids int32 [total], lens int32 [n]); lens[i] == -1 marks a text this tokenizer declined."""
        n = len(texts)
        # one blob + offsets (qre_wordpiece_encode_packed): n bytes objects and n pointers cost more
        # Python time than the C++ threads need to tokenise
        blob = "".join(texts).encode("utf-8")
        tl = np.fromiter(map(len, texts), np.int64, n)
        if n and int(tl.sum()) != len(blob):             # some text is not ASCII: byte lengths differ
            tl = np.fromiter((len(t.encode("utf-8")) for t in texts), np.int64, n)
        off = np.zeros(n + 1, np.int64)
        np.cumsum(tl, out=off[1:])
        cap = int(np.minimum(tl + 2, max_len).sum()) if n else 0
        ids = np.empty(max(cap, 1), np.int32)
        lens = np.empty(n, np.int32)
        _check(self._lib, self._lib.qre_wordpiece_encode_packed(self._h, C.c_int64(n), C.c_char_p(blob), _ptr(off),
                                                                C.c_int(max_len), _ptr(ids), C.c_int64(cap),
                                                                _ptr(lens), C.c_int(n_threads)))
        return ids[: int(lens[lens > 0].sum())], lens
