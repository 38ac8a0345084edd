"""Where does the host-visible latency of a small lookup+route call go?  Times (a) the Python
wrapper, (b) the bare C call with pre-built arguments, for batch 1 on a 1M x 384 cache."""
import argparse, ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]
import bench as B
from qre_b200 import _cabi
import torch

args = argparse.Namespace(entries=1_000_000, queries=4096, dim=384, centroids=8, devices=2, keep_rows=False)
torch.cuda.set_device(0)
wq = B.make_queries(args)
sh, _ = B.build_shard(args, wq, 0, 1, torch)
cfg = _cabi.RouteConfig(ttl_us=B.TTL_US)
q_ctx = wq["ctx_tab"][wq["q_ci"]]
m = 1
q, c, tok, hd, hc = wq["Q"][:m].copy(), q_ctx[:m].copy(), wq["tok"][:m].copy(), wq["heur_dev"][:m].copy(), wq["heur_conf"][:m].copy()
for _ in range(50):
    sh.lookup_route(q, c, cfg, B.NOW_US, tok=tok, heur_dev=hd, heur_conf=hc)
t = []
for _ in range(500):
    t0 = time.perf_counter(); sh.lookup_route(q, c, cfg, B.NOW_US, tok=tok, heur_dev=hd, heur_conf=hc); t.append(time.perf_counter() - t0)
print("python wrapper  p50 %.1f us" % (np.percentile(t, 50) * 1e6))
out = sh._host_outputs(m, True, 2, False)
res = _cabi.Result(**{f: _cabi._ptr(out.get(f)) for f in _cabi.RESULT_FIELDS})
p = cfg.to_c(B.NOW_US, True)
a = (sh.h, C.c_int(m), _cabi._ptr(q), _cabi._ptr(c), _cabi._ptr(tok), _cabi._ptr(hd), _cabi._ptr(hc), C.byref(p), C.byref(res))
fn = sh.lib.qre_lookup_route
t = []
for _ in range(500):
    t0 = time.perf_counter(); fn(*a); t.append(time.perf_counter() - t0)
print("bare C call     p50 %.1f us" % (np.percentile(t, 50) * 1e6))
p0 = cfg.to_c(B.NOW_US, False)
a0 = (sh.h, C.c_int(m), _cabi._ptr(q), _cabi._ptr(c), None, None, None, C.byref(p0), C.byref(res))
t = []
for _ in range(500):
    t0 = time.perf_counter(); fn(*a0); t.append(time.perf_counter() - t0)
print("bare C, lookup only p50 %.1f us" % (np.percentile(t, 50) * 1e6))
sh.profile(True)
for _ in range(100): fn(*a)
ms, n = sh.profile_read()
print("scan kernel %.1f us" % (ms / n * 1e3))
