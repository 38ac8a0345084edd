"""Token ids -> embedding (data-parallel over queries) -> all-gather -> row-sharded lookup + route, N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_fused_multi.py [--queries 4096] [--rows 1000000] [--steps 50]

Rank r embeds queries [r m/N, (r+1) m/N) and holds cache rows {i : i % N == r}; see
ShardedLookup.embed_ids / step.  Strong scaling: total work fixed.  Device-timed, max over ranks.
Random weights of the all-MiniLM-L6-v2 shape, synthetic ids and cache rows."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--queries", type=int, default=4096)
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--mean-len", type=float, default=24.0)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from qre_b200 import _cabi, synth
    from qre_b200.sharded import ShardedLookup
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    m, H = a.queries, 384
    rs = np.random.Generator(np.random.PCG64(2))
    lens = np.clip(np.round(rs.lognormal(np.log(a.mean_len) - 0.18, 0.6, m)), 3, 256).astype(np.int64)
    off = np.zeros(m + 1, np.int32)
    off[1:] = np.cumsum(lens)
    ids = rs.integers(0, 30522, int(off[-1])).astype(np.int32)
    per = (m + world - 1) // world
    enc = _cabi.Encoder(_cabi.EncoderConfig(), synth.encoder_random_weights(seed=1),
                        max_tokens=int(max(off[min(m, (r + 1) * per)] - off[min(m, r * per)] for r in range(world))) + 256,
                        max_seqs=per, device=local)
    n_local = (a.rows - rank + world - 1) // world
    sh = _cabi.Shard(dim=H, capacity=n_local, max_batch=m, device=local)
    rr = np.random.Generator(np.random.PCG64(100 + rank))
    for i0 in range(0, n_local, 1 << 17):
        k = min(1 << 17, n_local - i0)
        sh.append(rr.standard_normal((k, H), dtype=np.float32), np.full(k, 2, np.uint64), np.zeros(k, np.int64),
                  np.arange(i0 + 1, i0 + k + 1, dtype=np.uint64) * world + rank, (np.arange(k) % 2).astype(np.uint8),
                  np.full(k, 0.9))
    sh.set_centroids(np.random.Generator(np.random.PCG64(5)).standard_normal((8, H), dtype=np.float32),
                     (np.arange(8) % 2).astype(np.uint8))
    look = ShardedLookup(sh, _cabi.RouteConfig(ttl_us=10**15), rank, world, exchange="p2p" if world >= 4 else "nccl")
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)   # noqa: E731
    ctx_d, tok_d = t(np.full(m, 2, np.int64)), t(lens.astype(np.int32))
    hd_d, hc_d = t(np.zeros(m, np.uint8)), t(np.full(m, 0.8))

    def step():
        q = look.embed_ids(enc, ids, off)
        return look.step(q, ctx_d, 1, tok_d, hd_d, hc_d)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"metric": "token ids -> embedded -> looked up + routed, queries/s", "value": m / ms.item() * 1e3,
                          "unit": "queries/s", "n_gpus": world, "ms_per_batch": ms.item(), "queries": m,
                          "tokens": int(off[-1]), "cache_rows": a.rows, "scaling": "strong",
                          "exchange": look.exchange, "hits": int((out["hit_slot"] >= 0).sum().item())}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
