"""On-GPU replacement for the reference's sentence embedder.

The reference builds `SentenceTransformer("all-MiniLM-L6-v2")` (src/query_router_engine.py:126,
511) and calls `.encode(list_of_texts)` → float32 [n, 384] (qre:152-153, 176-177, 181, 571;
src/cache.py:412).  `B200SentenceEncoder` offers the same call on top of libqre_b200's encoder
(include/qre_encoder_b200.h): WordPiece on C++ host threads, the BERT encoder, mean pooling and
L2 normalisation on the GPU.  Pass an instance as `config["embedder"]` to the drop-in
`QueryRouter` / `SemanticRouter`; nothing else changes.

Weights and vocabulary are the caller's: this image has no network, so nothing is downloaded.
`from_hf_dir` reads a local Hugging Face model directory (config.json, vocab.txt,
model.safetensors or pytorch_model.bin) such as a sentence-transformers cache entry.

There is no CPU path: without the CUDA library or an sm_100 device the constructor raises.
"""
from __future__ import annotations

import json
import os
import threading
import unicodedata
from typing import Callable, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np

from . import _cabi

_HF_PREFIXES = ("bert.", "0.auto_model.", "auto_model.", "model.")


def _strip_prefix(sd: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    """Accept state dicts saved from BertModel, BertFor*, or a sentence-transformers module."""
    if "embeddings.word_embeddings.weight" in sd:
        return sd
    for p in _HF_PREFIXES:
        if p + "embeddings.word_embeddings.weight" in sd:
            return {k[len(p):]: v for k, v in sd.items() if k.startswith(p)}
    raise KeyError("state dict has no embeddings.word_embeddings.weight (BertModel naming expected)")


# ---- Python fallback tokenizer (the few texts the C++ tokenizer hands back: capital sigma, characters on
# ---- which the Python and Rust reference tokenizers disagree, unassigned code points) ----------------
def _is_punct(ch: str) -> bool:
    cp = ord(ch)
    return (33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126
            or unicodedata.category(ch).startswith("P"))


def _is_cjk(cp: int) -> bool:
    return any(lo <= cp <= hi for lo, hi in ((0x4E00, 0x9FFF), (0x3400, 0x4DBF), (0x20000, 0x2A6DF),
                                              (0x2A700, 0x2B73F), (0x2B740, 0x2B81F), (0x2B820, 0x2CEAF),
                                              (0xF900, 0xFAFF), (0x2F800, 0x2FA1F)))


def wordpiece_unicode(text: str, vocab: Dict[str, int], max_len: int, lower: bool = True) -> List[int]:
    """BERT BasicTokenizer + WordPiece for arbitrary Unicode, in Python (rare path)."""
    buf = []
    for ch in text:
        cp = ord(ch)
        if cp == 0 or cp == 0xFFFD:
            continue
        cat = unicodedata.category(ch)
        if ch in "\t\n\r" or cat == "Zs":
            buf.append(" ")
        elif cat.startswith("C"):
            continue
        elif _is_cjk(cp):
            buf.extend((" ", ch, " "))
        else:
            buf.append(ch)
    text = "".join(buf)
    if lower:
        text = "".join(c for c in unicodedata.normalize("NFD", text) if unicodedata.category(c) != "Mn").lower()
    unk = vocab["[UNK]"]
    ids: List[int] = []
    for tok in text.split():
        words, cur = [], ""
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
        for w in words:
            if len(w) > 100:
                ids.append(unk)
                continue
            start, pieces = 0, []
            while start < len(w):
                end, hit = len(w), None
                while start < end:
                    sub = w[start:end] if start == 0 else "##" + w[start:end]
                    if sub in vocab:
                        hit = vocab[sub]
                        break
                    end -= 1
                if hit is None:
                    pieces = [unk]
                    break
                pieces.append(hit)
                start = end
            ids.extend(pieces)
    ids = ids[: max(0, max_len - 2)]
    return [vocab["[CLS]"]] + ids + [vocab["[SEP]"]]


class B200SentenceEncoder:
    """`encode(texts)` like `sentence_transformers.SentenceTransformer`, computed on a B200."""

    def __init__(self, weights: Dict[str, np.ndarray], vocab_path: str,
                 config: Optional[_cabi.EncoderConfig] = None, device: int = 0, max_batch: int = 4096,
                 max_tokens: Optional[int] = None, do_lower_case: bool = True,
                 fallback_tokenizer: Optional[Callable[[str, int], Sequence[int]]] = None):
        w = _strip_prefix({k: np.asarray(v) for k, v in weights.items()})
        if config is None:
            V, H = w["embeddings.word_embeddings.weight"].shape
            layers = 1 + max(int(k.split(".")[2]) for k in w if k.startswith("encoder.layer."))
            F = w["encoder.layer.0.intermediate.dense.weight"].shape[0]
            P = w["embeddings.position_embeddings.weight"].shape[0]
            heads = H // 32 if H <= 384 else H // 64          # MiniLM: 32-wide heads; BERT-base family: 64
            config = _cabi.EncoderConfig(V, H, layers, heads, F, P, w["embeddings.token_type_embeddings.weight"].shape[0],
                                         1e-12, min(256, P))
        self.config = config
        self.max_batch = int(max_batch)
        self.max_seq_length = int(config.max_seq_len)
        self.max_tokens = int(max_tokens) if max_tokens else self.max_batch * 64
        self.max_tokens = max(self.max_tokens, self.max_seq_length)
        self._enc = _cabi.Encoder(config, w, max_tokens=self.max_tokens, max_seqs=self.max_batch, device=device)
        self._wp = _cabi.WordPiece(vocab_path, do_lower_case)
        self._lower = do_lower_case
        self._vocab: Optional[Dict[str, int]] = None
        self._vocab_path = vocab_path
        self._fallback = fallback_tokenizer
        self.device = device
        # The reference calls encode() from Flask worker threads (app.py:126, threaded=True); the device handle
        # has one set of activation buffers and one stream, so calls are serialised here, like Engine's lock.
        self._lock = threading.Lock()

    # ---- construction helpers ----------------------------------------------------------------
    @classmethod
    def from_hf_dir(cls, path: str, **kw) -> "B200SentenceEncoder":
        """A local Hugging Face / sentence-transformers model directory."""
        sub = os.path.join(path, "0_Transformer")
        root = sub if os.path.isdir(sub) else path
        with open(os.path.join(root, "config.json")) as f:
            c = json.load(f)
        st = os.path.join(root, "model.safetensors")
        if os.path.isfile(st):
            from safetensors.numpy import load_file
            sd = load_file(st)
        else:
            import torch
            sd = {k: v.float().numpy() for k, v in torch.load(os.path.join(root, "pytorch_model.bin"),
                                                             map_location="cpu").items()}
        max_seq = 256
        sb = os.path.join(path, "sentence_bert_config.json")
        if os.path.isfile(sb):
            with open(sb) as f:
                max_seq = int(json.load(f).get("max_seq_length", max_seq))
        cfg = _cabi.EncoderConfig(c["vocab_size"], c["hidden_size"], c["num_hidden_layers"], c["num_attention_heads"],
                                  c["intermediate_size"], c["max_position_embeddings"], c.get("type_vocab_size", 2),
                                  c.get("layer_norm_eps", 1e-12), min(max_seq, c["max_position_embeddings"]))
        tc = os.path.join(root, "tokenizer_config.json")      # the tokenizer's own setting decides the casing
        if "do_lower_case" not in kw and os.path.isfile(tc):
            with open(tc) as f:
                kw["do_lower_case"] = bool(json.load(f).get("do_lower_case", True))
        return cls(sd, os.path.join(root, "vocab.txt"), config=cfg, **kw)

    # ---- tokenisation ------------------------------------------------------------------------
    def _load_vocab(self) -> Dict[str, int]:
        if self._vocab is None:
            with open(self._vocab_path, encoding="utf-8") as f:
                self._vocab = {line.rstrip("\r\n"): i for i, line in enumerate(f)}
        return self._vocab

    def tokenize(self, texts: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
        """-> (ids int32 [total], offsets int32 [n + 1]); `[CLS] pieces [SEP]`, truncated."""
        ids, lens = self._wp.encode(list(texts), max_len=self.max_seq_length)
        if (lens < 0).any():                      # texts the C++ tokenizer handed back: Python path, spliced in order
            parts, o = [], 0
            for t, n in zip(texts, lens):
                if n >= 0:
                    parts.append(ids[o:o + n])
                    o += n
                elif self._fallback is not None:
                    parts.append(np.asarray(self._fallback(t, self.max_seq_length), np.int32))
                else:
                    parts.append(np.asarray(wordpiece_unicode(t, self._load_vocab(), self.max_seq_length, self._lower),
                                            np.int32))
            lens = np.fromiter((p.size for p in parts), np.int32, len(parts))
            ids = np.concatenate(parts) if parts else np.zeros(0, np.int32)
        off = np.zeros(len(texts) + 1, np.int32)
        np.cumsum(lens, out=off[1:])
        return ids, off

    # ---- SentenceTransformer surface ---------------------------------------------------------
    def get_sentence_embedding_dimension(self) -> int:
        return int(self.config.hidden)

    def _batches(self, off: np.ndarray) -> Iterable[Tuple[int, int]]:
        """Consecutive [i0, i1) ranges within the sequence and token capacities."""
        n, i0 = off.shape[0] - 1, 0
        while i0 < n:
            i1 = min(n, i0 + self.max_batch)
            while i1 > i0 + 1 and off[i1] - off[i0] > self.max_tokens:
                i1 -= 1
            yield i0, i1
            i0 = i1

    def encode(self, sentences: Union[str, Sequence[str]], batch_size: int = 0, normalize_embeddings: bool = True,
               convert_to_numpy: bool = True, **_ignored) -> np.ndarray:
        """float32 [n, hidden] (or [hidden] for a single string), as the reference's call sites use it.
        all-MiniLM-L6-v2 ends in a Normalize module, so its `encode` output is unit-norm whatever
        `normalize_embeddings` says; pass False only for models without that module."""
        single = isinstance(sentences, str)
        texts = [sentences] if single else list(sentences)
        out = np.empty((len(texts), self.config.hidden), np.float32)
        if texts:
            ids, off = self.tokenize(texts)              # host threads, no shared state: outside the lock
            with self._lock:
                for i0, i1 in self._batches(off):
                    out[i0:i1] = self._enc.encode_packed(ids[off[i0]:off[i1]], off[i0:i1 + 1] - off[i0],
                                                         normalize=normalize_embeddings)
        return out[0] if single else out

    @property
    def engine(self) -> "_cabi.Encoder":
        """The C-ABI encoder handle, for device-resident use: `engine.encode_dev(...)` writes fp32 rows where
        `qre_scan_dev` / `qre_cache_append_dev` read them, and `ShardedLookup.embed_ids(engine, ids, offsets)`
        embeds a batch data-parallel over the GPUs of a row-sharded cache."""
        return self._enc

    def close(self) -> None:
        self._enc.close()
        self._wp.close()
