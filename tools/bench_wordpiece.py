"""Host tokenisation throughput (no GPU): libqre_b200's C++ WordPiece against transformers' BertTokenizer
(tokenizers/Rust backend) on the same synthetic 30k-piece vocabulary and ASCII texts.  Checks that both
produce the same ids, then times them."""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-llm_b200")]


def main():
    from qre_b200 import _cabi
    rs = np.random.Generator(np.random.PCG64(0))
    letters = np.array(list("abcdefghijklmnopqrstuvwxyz"))
    words = sorted({"".join(rs.choice(letters, rs.integers(2, 9))) for _ in range(24000)})
    suffixes = sorted({"##" + "".join(rs.choice(letters, rs.integers(1, 4))) for _ in range(5000)})
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + list("abcdefghijklmnopqrstuvwxyz0123456789") + \
        ["##" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"] + list("!?,.'-()") + words + suffixes
    d = tempfile.mkdtemp()
    vp = os.path.join(d, "vocab.txt")
    with open(vp, "w") as f:
        f.write("\n".join(vocab) + "\n")
    n = 16384
    texts = []
    for _ in range(n):
        k = int(np.clip(rs.lognormal(2.6, 0.6), 2, 120))
        ws = [words[i] + ("ing" if rs.random() < 0.2 else "") for i in rs.integers(0, len(words), k)]
        texts.append(" ".join(ws).capitalize() + rs.choice(["?", ".", "!", ""]))
    wp = _cabi.WordPiece(vp)
    ids, lens = wp.encode(texts, max_len=256)
    out = {"texts": n, "ids": int(ids.size), "mean_len": float(lens.mean())}
    for threads in (1, 0):
        t0 = time.perf_counter()
        for _ in range(3):
            wp.encode(texts, max_len=256, n_threads=threads)
        dt = (time.perf_counter() - t0) / 3
        out[f"cpp_{'1_thread' if threads == 1 else 'all_threads'}_texts_per_s"] = n / dt
    try:
        from transformers import BertTokenizer
        tok = BertTokenizer(vp, do_lower_case=True)
        ref = tok(texts, truncation=True, max_length=256)["input_ids"]
        o, same = 0, True
        for r, ln in zip(ref, lens):
            same = same and ids[o:o + ln].tolist() == r
            o += ln
        out["same_ids_as_BertTokenizer"] = bool(same)
        t0 = time.perf_counter()
        tok(texts, truncation=True, max_length=256)
        out["hf_tokenizers_texts_per_s"] = n / (time.perf_counter() - t0)
    except Exception as e:      # transformers absent: report our side only
        out["hf"] = repr(e)
    out["host_cores"] = os.cpu_count()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
