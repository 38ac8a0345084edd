"""CPU restatement of the query embedder the reference calls upstream of the lookup
(SURVEY §8f-1).  TEST INFRASTRUCTURE ONLY: imported by tests/, never by the product.

What the reference does: `SentenceTransformer("all-MiniLM-L6-v2").encode([query])[0]`
(src/query_router_engine.py:126, 181, 511, 571; src/cache.py:412).  The algorithm lives in
third-party code that is NOT under /root/reference:

  * sentence-transformers 3.4.1 (requirements.txt:52): the model's `modules.json` chains
    Transformer(max_seq_length=256, do_lower_case) -> Pooling(mean over attention-masked
    tokens, sum / clamp(count, 1e-9)) -> Normalize (torch F.normalize, p=2, eps=1e-12);
  * transformers' `BertModel` (post-LayerNorm encoder, learned absolute positions, exact
    erf GELU, LayerNorm eps 1e-12, softmax(QK^T / sqrt(head_dim))) with the all-MiniLM-L6-v2
    shape: 6 layers, hidden 384, 12 heads, FFN 1536, vocab 30522, 512 positions;
  * BERT WordPiece tokenisation (BasicTokenizer + greedy longest-match-first pieces,
    `[CLS] ... [SEP]`, truncation to max_seq_length).

The published weights and vocab are not in this image and cannot be fetched, so parity is
pinned differently from the lookup path: tests/test_minilm_oracle_cpu.py checks this file
against `transformers.BertModel` / `transformers.BertTokenizer` (both importable here and on
the GPU box) with seeded random weights of the same shape and a synthetic vocab.  Against the
real checkpoint: parity unpinned.

Weights are a dict of float32 numpy arrays under the Hugging Face `BertModel.state_dict()`
names (e.g. "encoder.layer.0.attention.self.query.weight").
"""
from __future__ import annotations

import math
import unicodedata
from dataclasses import dataclass
from typing import Dict, List, Sequence

import numpy as np


@dataclass(frozen=True)
class EncoderConfig:
    vocab_size: int = 30522
    hidden: int = 384
    layers: int = 6
    heads: int = 12
    intermediate: int = 1536
    max_positions: int = 512
    type_vocab: int = 2
    ln_eps: float = 1e-12
    max_seq_len: int = 256        # sentence-transformers' max_seq_length for this model


MINILM_L6 = EncoderConfig()


def _ln(x: np.ndarray, g: np.ndarray, b: np.ndarray, eps: float) -> np.ndarray:
    mu = x.mean(-1, keepdims=True, dtype=np.float32)
    var = np.mean((x - mu) ** 2, -1, keepdims=True, dtype=np.float32)   # biased, as torch LayerNorm
    return ((x - mu) / np.sqrt(var + np.float32(eps))) * g + b


try:                                   # scipy is only a faster erf; the values are the same
    from scipy.special import erf as _erf
except Exception:                      # pragma: no cover
    _erf = np.vectorize(math.erf, otypes=[np.float64])


def _gelu(x: np.ndarray) -> np.ndarray:
    """Exact GELU (`hidden_act="gelu"`): x * 0.5 * (1 + erf(x / sqrt(2)))."""
    return (x * 0.5 * (1.0 + _erf(x.astype(np.float64) / math.sqrt(2.0)))).astype(np.float32)


def token_states(w: Dict[str, np.ndarray], ids: np.ndarray, cfg: EncoderConfig = MINILM_L6) -> np.ndarray:
    """Last hidden state [L, hidden] of one unpadded sequence (attention mask all ones)."""
    ids = np.asarray(ids, np.int64)
    L, H, nh = ids.shape[0], cfg.hidden, cfg.heads
    hd = H // nh
    f32 = np.float32
    x = (w["embeddings.word_embeddings.weight"][ids]
         + w["embeddings.position_embeddings.weight"][:L]
         + w["embeddings.token_type_embeddings.weight"][0]).astype(f32)
    x = _ln(x, w["embeddings.LayerNorm.weight"], w["embeddings.LayerNorm.bias"], cfg.ln_eps).astype(f32)
    for l in range(cfg.layers):
        p = f"encoder.layer.{l}."
        q = x @ w[p + "attention.self.query.weight"].T + w[p + "attention.self.query.bias"]
        k = x @ w[p + "attention.self.key.weight"].T + w[p + "attention.self.key.bias"]
        v = x @ w[p + "attention.self.value.weight"].T + w[p + "attention.self.value.bias"]
        q = q.reshape(L, nh, hd).transpose(1, 0, 2)
        k = k.reshape(L, nh, hd).transpose(1, 0, 2)
        v = v.reshape(L, nh, hd).transpose(1, 0, 2)
        s = (q @ k.transpose(0, 2, 1)) / f32(math.sqrt(hd))
        s = s - s.max(-1, keepdims=True)
        e = np.exp(s)
        a = e / e.sum(-1, keepdims=True)
        ctx = (a @ v).transpose(1, 0, 2).reshape(L, H).astype(f32)
        y = ctx @ w[p + "attention.output.dense.weight"].T + w[p + "attention.output.dense.bias"]
        x = _ln(y + x, w[p + "attention.output.LayerNorm.weight"], w[p + "attention.output.LayerNorm.bias"],
                cfg.ln_eps).astype(f32)
        h = _gelu(x @ w[p + "intermediate.dense.weight"].T + w[p + "intermediate.dense.bias"])
        y = h @ w[p + "output.dense.weight"].T + w[p + "output.dense.bias"]
        x = _ln(y + x, w[p + "output.LayerNorm.weight"], w[p + "output.LayerNorm.bias"], cfg.ln_eps).astype(f32)
    return x


def encode_ids(w: Dict[str, np.ndarray], seqs: Sequence[np.ndarray], cfg: EncoderConfig = MINILM_L6,
               normalize: bool = True) -> np.ndarray:
    """Sentence embeddings [m, hidden]: mean over tokens, then L2-normalise (Pooling + Normalize)."""
    out = np.zeros((len(seqs), cfg.hidden), np.float32)
    for i, ids in enumerate(seqs):
        ids = np.asarray(ids)[: cfg.max_seq_len]
        if ids.size == 0:
            continue
        x = token_states(w, ids, cfg)
        e = x.sum(0, dtype=np.float32) / np.float32(max(float(ids.size), 1e-9))
        if normalize:
            e = e / np.float32(max(float(np.sqrt(np.sum(e * e, dtype=np.float32))), 1e-12))
        out[i] = e
    return out


def random_weights(cfg: EncoderConfig = MINILM_L6, seed: int = 0, std: float = 0.05) -> Dict[str, np.ndarray]:
    """Seeded random weights of the model's shape (no checkpoint in this image).  LayerNorm gains
    near 1 and non-zero biases everywhere so that no term is silently dropped by a kernel."""
    rs = np.random.Generator(np.random.PCG64(seed))
    H, F = cfg.hidden, cfg.intermediate
    n = lambda *s: (rs.standard_normal(s) * std).astype(np.float32)   # noqa: E731
    g = lambda: (1.0 + 0.1 * rs.standard_normal(H)).astype(np.float32)   # noqa: E731
    w = {
        "embeddings.word_embeddings.weight": n(cfg.vocab_size, H),
        "embeddings.position_embeddings.weight": n(cfg.max_positions, H),
        "embeddings.token_type_embeddings.weight": n(cfg.type_vocab, H),
        "embeddings.LayerNorm.weight": g(),
        "embeddings.LayerNorm.bias": n(H),
    }
    for l in range(cfg.layers):
        p = f"encoder.layer.{l}."
        for nm in ("attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense"):
            w[p + nm + ".weight"] = n(H, H)
            w[p + nm + ".bias"] = n(H)
        w[p + "attention.output.LayerNorm.weight"] = g()
        w[p + "attention.output.LayerNorm.bias"] = n(H)
        w[p + "intermediate.dense.weight"] = n(F, H)
        w[p + "intermediate.dense.bias"] = n(F)
        w[p + "output.dense.weight"] = n(H, F)
        w[p + "output.dense.bias"] = n(H)
        w[p + "output.LayerNorm.weight"] = g()
        w[p + "output.LayerNorm.bias"] = n(H)
    return w


# ---------------------------------------------------------------------------------------------
# WordPiece (BERT uncased): BasicTokenizer + greedy longest-match-first.
# ---------------------------------------------------------------------------------------------
def _is_punct(ch: str) -> bool:
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith("P")


def _is_cjk(cp: int) -> bool:
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF
            or 0x2A700 <= cp <= 0x2B73F or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF
            or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


def basic_tokenize(text: str, lower: bool = True) -> List[str]:
    out = []
    for ch in text:                       # clean: drop NUL / U+FFFD / control, whitespace -> ' '
        cp = ord(ch)
        if cp == 0 or cp == 0xFFFD:
            continue
        cat = unicodedata.category(ch)
        if ch in "\t\n\r":
            out.append(" ")
        elif cat.startswith("C"):
            continue
        elif cat == "Zs":
            out.append(" ")
        elif _is_cjk(cp):
            out.append(f" {ch} ")
        else:
            out.append(ch)
    text = "".join(out)
    if lower:                             # BertNormalizer: strip accents (NFD, drop Mn), then lowercase
        text = "".join(c for c in unicodedata.normalize("NFD", text) if unicodedata.category(c) != "Mn")
        text = text.lower()
    words: List[str] = []
    for tok in text.split():
        cur = ""
        for ch in tok:
            if _is_punct(ch):
                if cur:
                    words.append(cur)
                    cur = ""
                words.append(ch)
            else:
                cur += ch
        if cur:
            words.append(cur)
    return words


def wordpiece(word: str, vocab: Dict[str, int], unk: str = "[UNK]", max_chars: int = 100) -> List[str]:
    if len(word) > max_chars:
        return [unk]
    pieces, start = [], 0
    while start < len(word):
        end, cur = len(word), None
        while start < end:
            sub = word[start:end]
            if start > 0:
                sub = "##" + sub
            if sub in vocab:
                cur = sub
                break
            end -= 1
        if cur is None:
            return [unk]
        pieces.append(cur)
        start = end
    return pieces


def tokenize_ids(text: str, vocab: Dict[str, int], max_seq_len: int = 256, lower: bool = True) -> np.ndarray:
    """`[CLS] pieces [SEP]` truncated to max_seq_len (longest-first on a single sequence = cut the tail)."""
    ids = []
    for wd in basic_tokenize(text, lower):
        ids.extend(vocab[p] for p in wordpiece(wd, vocab))
    ids = ids[: max(0, max_seq_len - 2)]
    return np.asarray([vocab["[CLS]"]] + ids + [vocab["[SEP]"]], np.int32)
