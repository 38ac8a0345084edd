"""How often does the bf16 on-GPU embedder flip a DECISION that fp32 embeddings of the same weights
would make?    Lives under tests/ because only tests may import oracle/.

    python tests/measure_decision_flips.py [--texts 1536]

Truth = the fp32 numpy restatement (oracle/minilm_oracle.py); candidate = the CUDA encoder (bf16
activations).  Random weights of the all-MiniLM-L6-v2 shape (no checkpoint offline) give embeddings
that are almost parallel, so the measured quantity is scale-free: the error of every pairwise
cosine, and the flip rate of `cos >= t` for thresholds t placed at quantiles of the actual cosine
distribution (where pairs are densest — the worst place for a threshold), plus the same for the
semantic router's margin rule |cos(q,c1) - cos(q,c0)| < 0.03 and its 0.05 relevance floor with
centroids = means of label embeddings (qre:140-178).  A flip can only happen when the fp32 value
lies within the embedding error of the threshold; the error bound printed here turns any pair
density into a flip bound for a real checkpoint."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--texts", type=int, default=1536)
    a = ap.parse_args()
    from oracle import minilm_oracle as MO
    from qre_b200 import _cabi, synth
    o = MO.EncoderConfig()
    w = synth.encoder_random_weights(seed=1)
    rs = np.random.Generator(np.random.PCG64(7))
    n = a.texts
    lens = np.clip(np.round(rs.lognormal(np.log(24.0) - 0.18, 0.6, n)), 3, 128).astype(np.int64)
    seqs = [rs.integers(0, o.vocab_size, k).astype(np.int32) for k in lens]
    enc = _cabi.Encoder(_cabi.EncoderConfig(), w, max_tokens=int(lens.sum()) + 256, max_seqs=n)
    got = enc.encode_ids(seqs).astype(np.float64)
    enc.close()
    want = np.concatenate([MO.encode_ids(w, seqs[i:i + 64], o) for i in range(0, n, 64)]).astype(np.float64)
    # centred copies: random-weight embeddings share a large common component; removing the mean
    # spreads the cosines over a realistic range, which is where thresholds bite
    res = {"texts": n, "pairs": n * (n - 1) // 2, "min_cosine_to_fp32": float((got * want).sum(1).min())}
    for tag, G, W in (("raw", got, want), ("centred", got - want.mean(0), want - want.mean(0))):
        Gn, Wn = G / np.linalg.norm(G, axis=1, keepdims=True), W / np.linalg.norm(W, axis=1, keepdims=True)
        iu = np.triu_indices(n, 1)
        cg, cw = (Gn @ Gn.T)[iu], (Wn @ Wn.T)[iu]
        err = np.abs(cg - cw)
        r = {"cos_fp32_quantiles": {q: float(np.quantile(cw, q)) for q in (0.01, 0.5, 0.99)},
             "abs_err_max": float(err.max()), "abs_err_p999": float(np.quantile(err, 0.999)),
             "abs_err_mean": float(err.mean()), "flip_rate_at_threshold": {}}
        for q in (0.5, 0.9, 0.99):
            t = float(np.quantile(cw, q))
            r["flip_rate_at_threshold"][f"q{q}"] = {"threshold": t, "flips": int(np.sum((cg >= t) != (cw >= t))),
                                                    "rate": float(np.mean((cg >= t) != (cw >= t)))}
        # semantic rules against two centroids built from the first 2 x 30 embeddings
        c0g, c1g, c0w, c1w = G[:30].mean(0), G[30:60].mean(0), W[:30].mean(0), W[30:60].mean(0)
        cosv = lambda X, c: (X @ c) / (np.linalg.norm(X, axis=1) * np.linalg.norm(c))   # noqa: E731
        mg, mw = np.abs(cosv(G, c1g) - cosv(G, c0g)), np.abs(cosv(W, c1w) - cosv(W, c0w))
        tg, tw = np.maximum(cosv(G, c1g), cosv(G, c0g)), np.maximum(cosv(W, c1w), cosv(W, c0w))
        r["semantic_rules"] = {
            "queries": n, "margin_abs_err_max": float(np.abs(mg - mw).max()),
            "ambiguity_rule_flips_0.03": int(np.sum((mg < 0.03) != (mw < 0.03))),
            "irrelevance_rule_flips_0.05": int(np.sum((tg < 0.05) != (tw < 0.05))),
            "device_flips_outside_band": int(np.sum(((cosv(G, c1g) > cosv(G, c0g)) != (cosv(W, c1w) > cosv(W, c0w)))
                                                    & (mw >= 0.03))),
            "fp32_margin_below_0.03": int(np.sum(mw < 0.03))}
        res[tag] = r
    print(json.dumps(res))


if __name__ == "__main__":
    main()
