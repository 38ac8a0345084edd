"""Throughput and accuracy of the on-GPU query embedder on one B200.

    python tools/bench_embed.py [--queries 4096] [--mean-len 24] [--steps 30] [--check 64]

Random weights of the all-MiniLM-L6-v2 shape (no checkpoint in this image), synthetic token ids,
ragged lengths.  Timed with CUDA events on the launching stream, inputs resident on the device.
Prints one JSON line: queries/s, tokens/s, algorithmic TFLOP/s against the measured bf16 peak, and
the error against the fp32 oracle on a sample (the tolerance the GPU tests enforce)."""
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
    ap.add_argument("--check", type=int, default=64, help="sequences compared with the fp32 oracle")
    a = ap.parse_args()
    import torch
    from oracle import minilm_oracle as MO
    from qre_b200 import _cabi
    cfg_o, cfg = MO.MINILM_L6, _cabi.EncoderConfig()
    w = MO.random_weights(cfg_o, seed=1)
    rs = np.random.Generator(np.random.PCG64(2))
    m = a.queries
    if a.fixed_len:
        lens = np.full(m, a.fixed_len, np.int64)
    else:   # log-normal around the mean, at least [CLS] x [SEP], at most max_seq_length
        lens = np.clip(np.round(rs.lognormal(np.log(a.mean_len) - 0.18, 0.6, m)), 3, cfg_o.max_seq_len).astype(np.int64)
    off = np.zeros(m + 1, np.int32)
    off[1:] = np.cumsum(lens)
    T = int(off[-1])
    ids = rs.integers(0, cfg_o.vocab_size, T).astype(np.int32)
    dev = torch.device("cuda", 0)
    enc = _cabi.Encoder(cfg, w, max_tokens=T, max_seqs=m)
    ids_d, off_d = torch.from_numpy(ids).to(dev), torch.from_numpy(off).to(dev)
    out_d = torch.empty((m, cfg_o.hidden), dtype=torch.float32, device=dev)
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
    H, F, Ls = cfg_o.hidden, cfg_o.intermediate, cfg_o.layers
    gemm_flops = 2.0 * T * Ls * (4 * H * H + 2 * H * F)
    attn_flops = 4.0 * float((lens.astype(np.float64) ** 2).sum()) * H * Ls
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    got = out_d.cpu().numpy()
    k = min(a.check, m)
    sel = np.argsort(lens)[np.linspace(0, m - 1, k).astype(int)]
    want = MO.encode_ids(w, [ids[off[i]:off[i + 1]] for i in sel], cfg_o)
    cos = (got[sel] * want).sum(1)
    line = {
        "metric": "embedded queries/s (all-MiniLM-L6-v2 shape, random weights)", "value": m / ms * 1e3, "unit": "queries/s",
        "tokens_per_s": T / ms * 1e3, "ms_per_batch": ms, "queries": m, "tokens": T, "mean_len": float(lens.mean()),
        "max_len": max_len, "launches_per_batch": (enc.launch_count() - l0) // a.steps,
        "algorithmic_tflops": (gemm_flops + attn_flops) / ms / 1e9, "gemm_share_of_flops": gemm_flops / (gemm_flops + attn_flops),
        "peaks": peaks, "dtype": "bf16 x bf16 -> fp32, bf16 activations",
        "vs_fp32_oracle": {"sequences": int(k), "min_cosine": float(cos.min()), "max_abs_diff": float(np.abs(got[sel] - want).max())},
    }
    print(json.dumps(line))
    enc.close()


if __name__ == "__main__":
    main()
