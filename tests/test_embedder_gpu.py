"""GPU: the SentenceTransformer-shaped host class over the on-GPU encoder (qre_b200/embedder.py) and
its use by the drop-in router through `config["embedder"]` (call sites qre:181, 571)."""
import os

import numpy as np
import pytest

from oracle import minilm_oracle as MO
from tests import replay_common as R
from tests.wordpiece_cases import TEXTS, VOCAB

pytestmark = pytest.mark.gpu

CFG_O = MO.EncoderConfig(vocab_size=len(VOCAB), hidden=128, layers=2, heads=4, intermediate=256, max_positions=64,
                         max_seq_len=48)


def _encoder(tmp_path, **kw):
    from qre_b200 import _cabi
    from qre_b200.embedder import B200SentenceEncoder
    vp = os.path.join(tmp_path, "vocab.txt")
    with open(vp, "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    w = MO.random_weights(CFG_O, seed=21)
    cfg = _cabi.EncoderConfig(len(VOCAB), 128, 2, 4, 256, 64, 2, 1e-12, 48)
    return B200SentenceEncoder(w, vp, config=cfg, **kw), w, {t: i for i, t in enumerate(VOCAB)}


def test_encode_matches_oracle_pipeline(tmp_path):
    enc, w, vocab = _encoder(tmp_path, max_batch=4, max_tokens=64)      # small capacities: exercises batching
    got = enc.encode(TEXTS)
    want = MO.encode_ids(w, [MO.tokenize_ids(t, vocab, 48) for t in TEXTS], CFG_O)
    assert got.shape == want.shape == (len(TEXTS), 128) and got.dtype == np.float32
    cos = (got * want).sum(1)
    assert cos.min() >= 0.999, float(cos.min())
    assert np.abs(got - want).max() <= 8e-3
    one = enc.encode("Hello worlds!")
    assert one.shape == (128,) and np.array_equal(one, enc.encode(["Hello worlds!"])[0])
    assert enc.encode([]).shape == (0, 128)
    assert enc.get_sentence_embedding_dimension() == 128 and enc.engine.launch_count() > 0
    ids, off = enc.tokenize(["hello world", "café 中文"])          # C++ path + Unicode fallback
    assert ids[off[0]:off[1]].tolist() == MO.tokenize_ids("hello world", vocab, 48).tolist()
    assert ids[off[1]:off[2]].tolist() == MO.tokenize_ids("café 中文", vocab, 48).tolist()
    enc.close()


def test_router_embeds_on_the_gpu(tmp_path):
    """QueryRouter(config["embedder"] = B200SentenceEncoder): misses, then exact-hash hits, then
    semantic hits for texts that tokenise identically (extra spaces / case) — the embedding the
    cache stored and the one the lookup computes come out of the same kernels, bit for bit."""
    enc, w, vocab = _encoder(tmp_path, max_batch=64)
    _, qre, _ = R.load_modules()
    cfg = dict(qre.PRODUCTION_CFG)
    cfg["embedder"] = enc
    cfg["centroids"] = enc.encode(["what is the world", "explain how to write pythonic code in python"])
    cfg["centroid_owner"] = [0, 1]
    router = qre.QueryRouter("hybrid", cfg)
    assert router.cache_embedder is enc
    words = ["hello", "world", "what", "is", "the", "explain", "how", "to", "write", "code", "python", "new", "york"]
    rs = np.random.Generator(np.random.PCG64(3))
    qs = sorted({" ".join(rs.choice(words, rs.integers(3, 9))) for _ in range(24)})
    first = router.route_queries(qs, context_keys=["k"] * len(qs))
    assert {d.method for d in first} == {"hybrid"} and not any(d.cache_hit for d in first)
    again = router.route_queries(qs, context_keys=["k"] * len(qs))
    assert {d.method for d in again} == {"hybrid_cached"}
    assert [d.device for d in again] == [d.device for d in first]
    respaced = [q.replace(" ", "   ") for q in qs]                       # different hash, same token ids
    sem = router.route_queries(respaced, context_keys=["k"] * len(qs))
    assert {d.method for d in sem} == {"hybrid_cached"}
    one = router.route_query(qs[0].replace(" ", "\t"), context_key="k")
    assert one.method == "hybrid_cached"
    st = router.get_cache_stats()
    assert st["hits"] >= 2 * len(qs) + 1
    enc.close()


def test_from_hf_dir_loads_a_saved_bert_checkpoint(tmp_path):
    """A BertModel saved by transformers (config.json + model.safetensors) plus vocab.txt and a
    sentence_bert_config.json, as a sentence-transformers cache entry has them: from_hf_dir must reproduce
    the torch pipeline (BertTokenizer -> BertModel -> mean pool -> normalise) on the same texts."""
    import json
    import torch
    from transformers import BertConfig, BertModel, BertTokenizer
    from qre_b200.embedder import B200SentenceEncoder
    d = os.path.join(tmp_path, "model")
    hc = BertConfig(vocab_size=len(VOCAB), hidden_size=128, num_hidden_layers=2, num_attention_heads=4,
                    intermediate_size=256, max_position_embeddings=64, hidden_act="gelu", layer_norm_eps=1e-12)
    torch.manual_seed(3)
    model = BertModel(hc, add_pooling_layer=False).eval()
    with torch.no_grad():                         # default init leaves LayerNorm at identity and biases at zero
        for n_, p_ in model.named_parameters():
            if "LayerNorm" in n_ or n_.endswith("bias"):
                p_.add_(0.1 * torch.randn_like(p_))
            else:
                p_.mul_(2.5)
    model.save_pretrained(d)
    with open(os.path.join(d, "vocab.txt"), "w", encoding="utf-8") as f:
        f.write("\n".join(VOCAB) + "\n")
    with open(os.path.join(d, "sentence_bert_config.json"), "w") as f:
        json.dump({"max_seq_length": 32, "do_lower_case": False}, f)
    enc = B200SentenceEncoder.from_hf_dir(d, max_batch=16)
    assert enc.max_seq_length == 32 and enc.config.heads == 4 and enc.config.layers == 2
    tok = BertTokenizer(os.path.join(d, "vocab.txt"), do_lower_case=True)
    texts = [t for t in TEXTS if t] + ["hello " * 50]
    b = tok(texts, truncation=True, max_length=32, padding=True, return_tensors="pt")
    with torch.no_grad():
        h = model(input_ids=b["input_ids"], attention_mask=b["attention_mask"]).last_hidden_state
    mf = b["attention_mask"].unsqueeze(-1).float()
    want = torch.nn.functional.normalize((h * mf).sum(1) / mf.sum(1).clamp(min=1e-9), p=2, dim=1).numpy()
    got = enc.encode(texts)
    cos = (got * want).sum(1)
    assert cos.min() >= 0.999, float(cos.min())
    assert np.abs(got - want).max() <= 8e-3
    enc.close()


def test_encode_from_several_threads(tmp_path):
    """The reference embeds from Flask worker threads; concurrent encode() calls on one encoder must give
    what serial calls give."""
    import threading
    enc, _, _ = _encoder(tmp_path, max_batch=8)
    texts = [t for t in TEXTS if t]
    serial = [enc.encode([t])[0] for t in texts]
    out, errs = {}, []

    def work(k):
        try:
            for rep in range(20):
                for i in range(k, len(texts), 4):
                    out[(k, rep, i)] = enc.encode([texts[i]])[0]
        except Exception as e:      # pragma: no cover
            errs.append(e)

    th = [threading.Thread(target=work, args=(k,)) for k in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs
    assert all(np.array_equal(v, serial[i]) for (_, _, i), v in out.items()) and len(out) == 20 * len(texts)
    enc.close()
