"""Observed error of the on-GPU embedder against the fp32 restatement (oracle/minilm_oracle.py) at a full
model shape; the GPU tests enforce the tolerance, this prints what was actually seen (recorded in
profiles/r1_embedder.md).  Lives under tests/ because only tests may import oracle/.

    python tests/measure_encoder_error.py [--shape minilm-l6|minilm-l12|bert-base] [--sequences 64]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="minilm-l6", choices=["minilm-l6", "minilm-l12", "bert-base"])
    ap.add_argument("--sequences", type=int, default=64)
    a = ap.parse_args()
    from oracle import minilm_oracle as MO
    from qre_b200 import _cabi, synth
    H, L, nh, F = {"minilm-l6": (384, 6, 12, 1536), "minilm-l12": (384, 12, 12, 1536), "bert-base": (768, 12, 12, 3072)}[a.shape]
    o = MO.EncoderConfig(hidden=H, layers=L, heads=nh, intermediate=F)
    w = synth.encoder_random_weights(hidden=H, layers=L, intermediate=F, seed=1)
    rs = np.random.Generator(np.random.PCG64(2))
    lens = np.clip(np.round(rs.lognormal(np.log(24.0) - 0.18, 0.6, 4096)), 3, o.max_seq_len).astype(np.int64)
    lens = np.sort(lens)[np.linspace(0, 4095, a.sequences).astype(int)]          # the whole length range
    seqs = [rs.integers(0, o.vocab_size, n).astype(np.int32) for n in lens]
    enc = _cabi.Encoder(_cabi.EncoderConfig(hidden=H, layers=L, heads=nh, intermediate=F), w,
                        max_tokens=int(lens.sum()) + 256, max_seqs=a.sequences)
    got = enc.encode_ids(seqs)
    want = MO.encode_ids(w, seqs, o)
    cos = (got * want).sum(1)
    print(json.dumps({"shape": a.shape, "sequences": a.sequences, "min_len": int(lens.min()), "max_len": int(lens.max()),
                      "min_cosine": float(cos.min()), "max_abs_diff": float(np.abs(got - want).max())}))
    enc.close()


if __name__ == "__main__":
    main()
