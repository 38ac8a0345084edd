"""TEST-ONLY stand-in for qre_b200.engine.Engine backed by the CPU oracle.

It lets the CPU suite exercise the HOST logic of the drop-in classes (exact-hash path, LRU
and TTL bookkeeping, eviction order, control flow, reasoning strings) without a GPU.  The
product never imports this: `cache.QueryCache` / `query_router_engine.*` construct the real
Engine, which fails loudly when libqre_b200.so or the GPU is missing."""
import numpy as np

from oracle import qre_oracle as O


class OracleEngine:
    def __init__(self, capacity=512, device=0, max_batch=64, dim=384, scan_mode=0):
        self.capacity = capacity
        self.dim = dim
        self.max_batch = max_batch
        self.rows = np.zeros((capacity, dim), np.float32)
        self.ctx = np.full(capacity, O.CTX_TOMBSTONE, np.uint64)
        self.ts = np.zeros(capacity, np.int64)
        self.lru = np.zeros(capacity, np.uint64)
        self.pdev = np.zeros(capacity, np.uint8)
        self.pconf = np.zeros(capacity, np.float64)
        self.hwm, self.free = 0, []
        self._centroids = self._owner = None
        self.calls = 0

    @property
    def shard(self):
        """Truthy once rows exist (the real Engine creates its device handle lazily; callers only test for None)."""
        return self if self.hwm else None

    def close(self):
        self.clear()

    def set_centroids(self, c, o):
        self._centroids, self._owner = np.asarray(c, np.float32), np.asarray(o, np.uint8)

    def set_perf_health(self, h):
        self._health = np.asarray(h, np.float64)

    @property
    def n_centroids(self):
        return 0 if self._centroids is None else self._centroids.shape[0]

    def _grow(self, need):
        while need > len(self.ctx):
            for name in ("rows", "ctx", "ts", "lru", "pdev", "pconf"):
                a = getattr(self, name)
                pad = np.zeros_like(a)
                if name == "ctx":
                    pad[:] = O.CTX_TOMBSTONE
                setattr(self, name, np.concatenate([a, pad]))
            self.capacity = len(self.ctx)

    def append(self, row, ctx, ts_us, lru, pdev, pconf):
        row = np.asarray(row, np.float32)
        if row.shape[0] != self.rows.shape[1]:          # dimension is learnt from the first row
            assert self.hwm == 0
            self.dim = row.shape[0]
            self.rows = np.zeros((len(self.ctx), self.dim), np.float32)
        if self.free:
            s = self.free.pop()
        else:
            self._grow(self.hwm + 1)
            s = self.hwm
            self.hwm += 1
        self.rows[s], self.ctx[s], self.ts[s], self.lru[s] = row, ctx, ts_us, lru
        self.pdev[s], self.pconf[s] = pdev, pconf
        return s

    def update(self, slot, row=None, ts_us=None, lru=None, pdev=None, pconf=None):
        if row is not None: self.rows[slot] = row
        if ts_us is not None: self.ts[slot] = ts_us
        if lru is not None: self.lru[slot] = lru
        if pdev is not None: self.pdev[slot] = pdev
        if pconf is not None: self.pconf[slot] = pconf

    def append_many(self, rows, ctx, ts_us, lru, pdev, pconf):
        return np.array([self.append(*a) for a in zip(rows, ctx, ts_us, lru, pdev, pconf)], np.int64)

    def update_many(self, slots, rows=None, ts_us=None, lru=None, pdev=None, pconf=None):
        for i, s_ in enumerate(slots):
            self.update(int(s_), None if rows is None else rows[i], None if ts_us is None else ts_us[i],
                        None if lru is None else lru[i], None if pdev is None else pdev[i],
                        None if pconf is None else pconf[i])

    def tombstone(self, slots):
        for s in slots:
            self.ctx[s] = O.CTX_TOMBSTONE
            self.free.append(int(s))

    def clear(self):
        self.ctx[:] = O.CTX_TOMBSTONE
        self.hwm, self.free = 0, []

    def read_rows(self, slots):
        return self.rows[list(slots)].copy()

    def scan(self, q, q_ctx, now_us, cfg, route_in=None):
        self.calls += 1
        q = np.ascontiguousarray(q, np.float32)
        if q.ndim == 1:
            q = q[None, :]
        n = self.hwm
        img = O.CacheImage(self.rows[:n], self.ctx[:n], self.ts[:n], self.lru[:n])
        m = q.shape[0]
        out = {}
        if n:
            rows, sims = zip(*[O.scan_faithful(q[i], int(q_ctx[i]), img, now_us, cfg.ttl_us,
                                               cfg.similarity_threshold) for i in range(m)])
        else:
            rows, sims = [-1] * m, [0.0] * m
        out["hit_slot"] = np.array(rows, np.int64)
        out["hit_sim"] = np.array(sims, np.float32)
        if route_in is not None:
            p = O.RouteParams(n_devices=cfg.n_devices, token_threshold=cfg.token_threshold,
                              w_token=cfg.w_token, w_semantic=cfg.w_semantic, w_heuristic=cfg.w_heuristic,
                              w_perf=cfg.w_perf,
                              sem=O.SemanticParams(cfg.sem_min_similarity, cfg.sem_margin_threshold))
            sims_c = np.array([[O.centroid_similarity(q[i], c) for c in self._centroids] for i in range(m)],
                              np.float32)
            r = O.route_batch(q, route_in["tok"], route_in["heur_dev"], route_in["heur_conf"],
                              self._centroids, self._owner, p)
            # use the reference's scalar cosine (bit-identical to the golden) for the decisions
            for i in range(m):
                td, tc = O.token_decision(int(route_in["tok"][i]), p.token_threshold)
                sd, sc, sm, _ = O.semantic_decision([float(x) for x in sims_c[i]], self._owner, p.n_devices,
                                                    td, tc, p.sem)
                d, c, S = O.hybrid_decision([(td, tc, p.w_token), (sd, sc, p.w_semantic),
                                             (int(route_in["heur_dev"][i]), float(route_in["heur_conf"][i]),
                                              p.w_heuristic)], p.n_devices)
                r["sem_dev"][i], r["sem_conf"][i], r["sem_method"][i] = sd, sc, sm
                r["dev"][i], r["conf"][i], r["S"][i] = d, c, S
            out.update(centroid_sim=sims_c, tok_dev=r["tok_dev"], tok_conf=r["tok_conf"],
                       sem_dev=r["sem_dev"], sem_conf=r["sem_conf"], sem_method=r["sem_method"],
                       dev=r["dev"], conf=r["conf"], score=r["S"])
        return out
