"""Throughput and accuracy of the on-GPU query embedder on one B200.

    python tools/bench_embed.py [--queries 4096] [--mean-len 24] [--steps 30] [--check 64]

Random weights of the all-MiniLM-L6-v2 shape (no checkpoint in this image), synthetic token ids,
ragged lengths.  Timed with CUDA events on the launching stream, inputs resident on the device.
Prints one JSON line: queries/s, tokens/s, algorithmic TFLOP/s against the measured bf16 peak.  The error
against the fp32 restatement is measured by tests/measure_encoder_error.py (tools do not import oracle/)."""
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
    ap.add_argument("--mean-len", type=float, default=24.0)
    ap.add_argument("--fixed-len", type=int, default=0)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--check", type=int, default=0, help="ignored (kept for old command lines)")
    ap.add_argument("--shape", default="minilm-l6", choices=["minilm-l6", "minilm-l12", "bert-base"],
                    help="encoder shape: all-MiniLM-L6-v2 (default), all-MiniLM-L12-v2, or BERT-base widths (768 / 12 x 64 / 3072)")
    ap.add_argument("--fused-cache-rows", type=int, default=0,
                    help="also time token ids -> embedding -> cache lookup + hybrid route against a cache of "
                         "this many random rows, everything on the device (SURVEY 8f-1 feeding 8a)")
    a = ap.parse_args()
    import torch
    from qre_b200 import _cabi, synth
    shapes = {"minilm-l6": (384, 6, 12, 1536), "minilm-l12": (384, 12, 12, 1536), "bert-base": (768, 12, 12, 3072)}
    Hh, Ll, Nh, Ff = shapes[a.shape]
    cfg = _cabi.EncoderConfig(hidden=Hh, layers=Ll, heads=Nh, intermediate=Ff)
    w = synth.encoder_random_weights(hidden=Hh, layers=Ll, intermediate=Ff, seed=1)
    rs = np.random.Generator(np.random.PCG64(2))
    m = a.queries
    if a.fixed_len:
        lens = np.full(m, a.fixed_len, np.int64)
    else:   # log-normal around the mean, at least [CLS] x [SEP], at most max_seq_length
        lens = np.clip(np.round(rs.lognormal(np.log(a.mean_len) - 0.18, 0.6, m)), 3, cfg.max_seq_len).astype(np.int64)
    off = np.zeros(m + 1, np.int32)
    off[1:] = np.cumsum(lens)
    T = int(off[-1])
    ids = rs.integers(0, cfg.vocab_size, T).astype(np.int32)
    dev = torch.device("cuda", 0)
    enc = _cabi.Encoder(cfg, w, max_tokens=T, max_seqs=m)
    ids_d, off_d = torch.from_numpy(ids).to(dev), torch.from_numpy(off).to(dev)
    out_d = torch.empty((m, cfg.hidden), dtype=torch.float32, device=dev)
    side = torch.cuda.Stream(device=dev)
    max_len = int(lens.max())

    def step():
        enc.encode_dev(m, ids_d.data_ptr(), off_d.data_ptr(), T, max_len, out_d.data_ptr(), stream=side.cuda_stream)

    with torch.cuda.stream(side):
        for _ in range(a.warmup):
            step()
        side.synchronize()
        l0 = enc.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(a.steps):
            step()
        e1.record(side)
        side.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    H, F, Ls = cfg.hidden, cfg.intermediate, cfg.layers
    gemm_flops = 2.0 * T * Ls * (4 * H * H + 2 * H * F)
    attn_flops = 4.0 * float((lens.astype(np.float64) ** 2).sum()) * H * Ls
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    got = out_d.cpu().numpy()
    line = {
        "metric": f"embedded queries/s ({a.shape} shape, random weights)", "value": m / ms * 1e3, "unit": "queries/s",
        "tokens_per_s": T / ms * 1e3, "ms_per_batch": ms, "queries": m, "tokens": T, "mean_len": float(lens.mean()),
        "max_len": max_len, "launches_per_batch": (enc.launch_count() - l0) // a.steps,
        "algorithmic_tflops": (gemm_flops + attn_flops) / ms / 1e9, "gemm_share_of_flops": gemm_flops / (gemm_flops + attn_flops),
        "peaks": peaks, "dtype": "bf16 x bf16 -> fp32, bf16 activations",
        "unit_norm_max_dev": float(np.abs(np.linalg.norm(got, axis=1) - 1.0).max()),
    }
    if a.fused_cache_rows > 0:
        from qre_b200.sharded import ShardedLookup
        n = a.fused_cache_rows
        sh = _cabi.Shard(dim=cfg.hidden, capacity=n + 256, max_batch=m)
        rr = np.random.Generator(np.random.PCG64(7))
        for i0 in range(0, n, 1 << 17):
            k = min(1 << 17, n - i0)
            rows = rr.standard_normal((k, cfg.hidden), dtype=np.float32)
            sh.append(rows, np.full(k, 2, np.uint64), np.zeros(k, np.int64), np.arange(i0 + 1, i0 + k + 1, dtype=np.uint64),
                      (np.arange(k) % 2).astype(np.uint8), np.full(k, 0.9))
        # Random-weight encoders map every text close to one direction (pairwise cosines > 0.85), unlike the
        # trained model, so each planted embedding is a hit for EVERY query: plant only 4 (a window of at most
        # 4 candidates per query; more than 8 would push every query through the overflow path).
        sh.append(got[:4], np.full(4, 2, np.uint64), np.zeros(4, np.int64),
                  np.arange(n + 1, n + 5, dtype=np.uint64), np.zeros(4, np.uint8), np.full(4, 0.9))
        sh.set_centroids(got[:8].copy(), (np.arange(8) % 2).astype(np.uint8))
        look = ShardedLookup(sh, _cabi.RouteConfig(ttl_us=10**15))
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)   # noqa: E731
        ctx_d, tok_d = t(np.full(m, 2, np.int64)), t(lens.astype(np.int32))
        hd_d, hc_d = t(np.zeros(m, np.uint8)), t(np.full(m, 0.8))
        q_d = torch.empty((m, cfg.hidden), dtype=torch.float32, device=dev)

        def fused():
            enc.encode_dev(m, ids_d.data_ptr(), off_d.data_ptr(), T, max_len, q_d.data_ptr(), stream=side.cuda_stream)
            return look.step(q_d, ctx_d, 1, tok_d, hd_d, hc_d)

        with torch.cuda.stream(side):
            for _ in range(a.warmup):
                out = fused()
            side.synchronize()
            e0.record(side)
            for _ in range(a.steps):
                out = fused()
            e1.record(side)
            side.synchronize()
        fms = e0.elapsed_time(e1) / a.steps
        hits = int((out["hit_slot"] >= 0).sum().item())
        print(json.dumps({"metric": "token ids -> embedded -> looked up + routed, queries/s (one B200, on device)",
                          "value": m / fms * 1e3, "unit": "queries/s", "ms_per_batch": fms, "queries": m, "tokens": T,
                          "cache_rows": n, "hits": hits, "embed_ms": ms, "lookup_route_ms": fms - ms}))
        sh.close()
    print(json.dumps(line))
    enc.close()


if __name__ == "__main__":
    main()
