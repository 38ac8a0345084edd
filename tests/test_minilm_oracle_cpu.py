"""Pins oracle/minilm_oracle.py (the embedder restatement) against the third-party
code the reference actually calls through sentence-transformers: `transformers.BertModel` and
`transformers.BertTokenizer`, with seeded random weights of the all-MiniLM-L6-v2 shape and a
synthetic vocab (the published checkpoint is not in this image)."""
import os

import numpy as np
import pytest

from oracle import minilm_oracle as MO

torch = pytest.importorskip("torch")
transformers = pytest.importorskip("transformers")

SMALL = MO.EncoderConfig(vocab_size=500, hidden=128, layers=2, heads=4, intermediate=256, max_positions=64,
                         max_seq_len=48)


def _hf_model(cfg: MO.EncoderConfig, w):
    from transformers import BertConfig, BertModel
    hc = BertConfig(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden, num_hidden_layers=cfg.layers,
                    num_attention_heads=cfg.heads, intermediate_size=cfg.intermediate,
                    max_position_embeddings=cfg.max_positions, type_vocab_size=cfg.type_vocab,
                    hidden_act="gelu", layer_norm_eps=cfg.ln_eps, hidden_dropout_prob=0.0,
                    attention_probs_dropout_prob=0.0)
    m = BertModel(hc, add_pooling_layer=False).eval()
    sd = {k: torch.from_numpy(v) for k, v in w.items()}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("position_ids" in k or "token_type_ids" in k for k in missing), missing
    return m


def _st_pipeline(model, seqs):
    """sentence-transformers' Transformer -> Pooling(mean) -> Normalize on a padded batch."""
    L = max(len(s) for s in seqs)
    ids = torch.zeros((len(seqs), L), dtype=torch.long)
    mask = torch.zeros((len(seqs), L), dtype=torch.long)
    for i, s in enumerate(seqs):
        ids[i, : len(s)] = torch.from_numpy(np.asarray(s, np.int64))
        mask[i, : len(s)] = 1
    with torch.no_grad():
        tok = model(input_ids=ids, attention_mask=mask).last_hidden_state
    mf = mask.unsqueeze(-1).float()
    emb = (tok * mf).sum(1) / mf.sum(1).clamp(min=1e-9)
    return torch.nn.functional.normalize(emb, p=2, dim=1).numpy()


WIDE_HEADS = MO.EncoderConfig(vocab_size=400, hidden=128, layers=2, heads=2, intermediate=320, max_positions=40,
                              max_seq_len=40)      # 64-wide heads, as the BERT-base family


@pytest.mark.parametrize("cfg,seed", [(SMALL, 3), (WIDE_HEADS, 4), (MO.MINILM_L6, 5)])
def test_encoder_restatement_equals_bertmodel(cfg, seed):
    w = MO.random_weights(cfg, seed=seed)
    model = _hf_model(cfg, w)
    rs = np.random.Generator(np.random.PCG64(seed))
    lens = [3, 11, 40] if cfg is MO.MINILM_L6 else [1, 2, 7, 16, 33, cfg.max_seq_len]
    seqs = [rs.integers(0, cfg.vocab_size, n).astype(np.int32) for n in lens]
    want = _st_pipeline(model, seqs)
    got = MO.encode_ids(w, seqs, cfg)
    assert got.shape == want.shape and got.dtype == np.float32
    # fp32 on both sides, different summation order only
    np.testing.assert_allclose(got, want, rtol=0, atol=2e-5)
    cos = (got * want).sum(1)
    assert cos.min() > 1 - 1e-6


def test_truncation_and_empty():
    w = MO.random_weights(SMALL, seed=1)
    long = np.arange(100, dtype=np.int32) % SMALL.vocab_size
    a = MO.encode_ids(w, [long], SMALL)
    b = MO.encode_ids(w, [long[: SMALL.max_seq_len]], SMALL)
    assert np.array_equal(a, b)
    z = MO.encode_ids(w, [np.zeros(0, np.int32)], SMALL)
    assert not z.any()


from tests.wordpiece_cases import TEXTS, VOCAB  # noqa: E402


def test_wordpiece_restatement_equals_berttokenizer(tmp_path):
    from transformers import BertTokenizer
    vp = os.path.join(tmp_path, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    tok = BertTokenizer(vp, do_lower_case=True)
    vocab = {t: i for i, t in enumerate(VOCAB)}
    for text in TEXTS:
        for max_len in (256, 16):
            want = tok(text, truncation=True, max_length=max_len)["input_ids"]
            got = MO.tokenize_ids(text, vocab, max_len).tolist()
            assert got == want, (text, max_len, got, want)


def test_wordpiece_restatement_on_random_texts(tmp_path):
    """1,500 generated texts (vocab words, ASCII punctuation and control characters, accents, CJK, emoji,
    odd whitespace): the restatement and BertTokenizer must produce the same ids for each."""
    import random
    from transformers import BertTokenizer
    vp = os.path.join(tmp_path, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    tok = BertTokenizer(vp, do_lower_case=True)
    vocab = {t: i for i, t in enumerate(VOCAB)}
    rnd = random.Random(11)
    words = [v for v in VOCAB if not v.startswith("[") and not v.startswith("##")]
    extra = ["É", "ü", "ñ", "ß", "ø", "中", "文", "日本", "😀", " ", " ", "​", "\x07", "\x7f", "…", "—", "«", "»",
             "ǅ", "İ", "ﬁ", "²", "½"]
    glue = [" ", "  ", "\t", "\n", "\r\n", "", "-", "'", ",", ".", "!", "?", "(", ")", "$", "_", "/", "\\", "[", "]", "#", "##"]
    for i in range(1500):
        parts = []
        for _ in range(rnd.randint(0, 14)):
            r = rnd.random()
            w = rnd.choice(words) if r < 0.6 else (rnd.choice(extra) if r < 0.8 else
                                                    "".join(rnd.choice("abcxyzHELLO019") for _ in range(rnd.randint(1, 7))))
            parts.append(w.upper() if rnd.random() < 0.15 else w)
            parts.append(rnd.choice(glue))
        text = "".join(parts)
        for max_len in (256, 9):
            want = tok(text, truncation=True, max_length=max_len)["input_ids"]
            got = MO.tokenize_ids(text, vocab, max_len).tolist()
            assert got == want, (i, repr(text), max_len, got, want)


def test_cpp_wordpiece_equals_berttokenizer_on_unicode_texts(tmp_path):
    """The product's C++ tokenizer against transformers' BertTokenizer ITSELF (backed by the Rust
    `tokenizers` package, what sentence-transformers loads for the reference's model) on 4,000 generated
    Unicode texts incl. arbitrary assigned code points, both casings: every text the C++ path accepts
    must give BertTokenizer's ids; the ones it hands back (-1) are those holding a character its tables
    list as ambiguous between the implementations."""
    from transformers import BertTokenizer
    from qre_b200 import _cabi
    from tests.test_wordpiece_cpu import UNI_VOCAB, unicode_texts
    vp = os.path.join(tmp_path, "vocab_u.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(UNI_VOCAB) + "\n")
    texts, n_mixed = unicode_texts(seed=29)
    for lower in (True, False):
        tok = BertTokenizer(vp, do_lower_case=lower)
        wp = _cabi.WordPiece(vp, do_lower_case=lower)
        ids, lens = wp.encode(texts, max_len=48)
        ref = tok(texts, truncation=True, max_length=48)["input_ids"]
        o, accepted = 0, 0
        for t, n, r in zip(texts, lens.tolist(), ref):
            if n < 0:
                continue
            assert ids[o:o + n].tolist() == r, (repr(t), lower)
            o += n
            accepted += 1
        assert accepted >= 0.93 * len(texts)
