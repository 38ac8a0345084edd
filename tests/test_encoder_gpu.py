"""GPU parity of the on-GPU query embedder (include/qre_encoder_b200.h) against
oracle/minilm_oracle.py — the fp32 restatement of what SentenceTransformer.encode computes
(src/query_router_engine.py:181, 571), itself pinned against transformers' BertModel.

Floating-point tolerance, stated here: the CUDA path multiplies in bf16 (fp32 accumulate) and keeps
activations in bf16 between kernels, the oracle is fp32 throughout.  For unit-norm embeddings we
require cosine >= 0.999 to the oracle and max |diff| <= 8e-3 per component (observed values are
recorded by tools/bench_embed.py); the GEMM hook alone is held to bf16 output rounding (2^-8
relative) plus accumulation-order slack."""
import numpy as np
import pytest

from oracle import minilm_oracle as MO

pytestmark = pytest.mark.gpu

COS_MIN = 0.999
ABS_MAX = 8e-3


def _bf16(x):
    """Round-to-nearest-even to bf16, returned as float32 (what the kernels' inputs/outputs are)."""
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def _gelu(x):
    return MO._gelu(x.astype(np.float32))


@pytest.mark.parametrize("rows,n,k", [(1, 192, 64), (100, 384, 384), (128, 1152, 384), (300, 1536, 384),
                                      (257, 384, 1536), (129, 96, 128), (500, 64, 64), (64, 800, 256),
                                      (40, 72, 64), (20000, 384, 384)])
@pytest.mark.parametrize("epi", [0, 1, 2])
def test_gemm_hook_matches_numpy(rows, n, k, epi):
    from qre_b200 import _cabi
    rs = np.random.Generator(np.random.PCG64(rows * 7 + n + k + epi))
    a = _bf16(rs.standard_normal((rows, k)).astype(np.float32))
    w = _bf16((rs.standard_normal((n, k)) / np.sqrt(k)).astype(np.float32))
    bias = rs.standard_normal(n).astype(np.float32)
    res = _bf16(rs.standard_normal((rows, n)).astype(np.float32)) if epi == 2 else None
    got = _cabi.encoder_debug_gemm(a, w, bias, res, epi)
    ref = a.astype(np.float64) @ w.astype(np.float64).T + bias
    if epi == 1:
        ref = _gelu(ref).astype(np.float64)
    elif epi == 2:
        ref = ref + res
    # bf16 output rounding (2^-9 relative after RNE) + fp32 accumulation slack
    tol = np.abs(ref) * 2.0 ** -8 + 2e-5 * np.sqrt(k)
    bad = np.abs(got - ref) > tol
    assert not bad.any(), (int(bad.sum()), float(np.abs(got - ref).max()))


@pytest.mark.parametrize("rows,k", [(1, 384), (300, 384), (1000, 1536), (20000, 384)])
def test_gemm_hook_layernorm_epilogue(rows, k):
    """Epilogue 3: LayerNorm(x W^T + bias + residual) over 384 features inside the projection kernel."""
    from qre_b200 import _cabi
    rs = np.random.Generator(np.random.PCG64(rows + k))
    n = 384
    a = _bf16(rs.standard_normal((rows, k)).astype(np.float32))
    w = _bf16((rs.standard_normal((n, k)) / np.sqrt(k)).astype(np.float32))
    bias = rs.standard_normal(n).astype(np.float32)
    res = _bf16((2.0 * rs.standard_normal((rows, n)) + 0.5).astype(np.float32))
    g = (1.0 + 0.2 * rs.standard_normal(n)).astype(np.float32)
    b = rs.standard_normal(n).astype(np.float32)
    got = _cabi.encoder_debug_gemm(a, w, bias, res, 3, ln_g=g, ln_b=b, ln_eps=1e-12)
    y = a.astype(np.float64) @ w.astype(np.float64).T + bias + res
    mu, sd = y.mean(1, keepdims=True), y.std(1, keepdims=True)
    ref = (y - mu) / sd * g + b
    # the kernel rounds y to bf16 before normalising (error 2^-9 |y|, i.e. 2^-9 (|ref - b| + |g mu| / sd) after
    # the scaling) and rounds the result to bf16 again (2^-9 |ref|); factor 2 of slack on both
    tol = 2.0 ** -8 * np.abs(ref) + 2.0 ** -8 * (np.abs(ref - b) + np.abs(g) * np.abs(mu) / sd) + 2e-3
    bad = np.abs(got - ref) > tol
    assert not bad.any(), (int(bad.sum()), float(np.abs(got - ref).max()))


def _check(cfg_o, ccfg, lens, seed, normalize=True):
    from qre_b200 import _cabi
    w = MO.random_weights(cfg_o, seed=seed)
    rs = np.random.Generator(np.random.PCG64(seed + 100))
    seqs = [rs.integers(0, cfg_o.vocab_size, n).astype(np.int32) for n in lens]
    enc = _cabi.Encoder(ccfg, w, max_tokens=max(128, int(sum(min(n, cfg_o.max_seq_len) for n in lens))),
                        max_seqs=len(lens))
    got = enc.encode_ids(seqs, normalize=normalize)
    want = MO.encode_ids(w, seqs, cfg_o, normalize=normalize)
    assert enc.launch_count() > 0
    enc.close()
    return got, want


def _assert_close(got, want, lens):
    nz = np.asarray(lens) > 0
    assert not got[~nz].any() and not want[~nz].any()
    cos = (got[nz] * want[nz]).sum(1) / (np.linalg.norm(got[nz], axis=1) * np.linalg.norm(want[nz], axis=1))
    assert cos.min() >= COS_MIN, float(cos.min())
    assert np.abs(got - want).max() <= ABS_MAX, float(np.abs(got - want).max())
    assert np.allclose(np.linalg.norm(got[nz], axis=1), 1.0, atol=1e-5)


def test_against_transformers_fixture(golden_dir):
    """tests/golden/embedder.npz holds what transformers' BertModel + mean pooling + normalisation returned for
    seeded weights and ids (oracle/gen_golden.py embedder): the CUDA path against the third-party code's own
    output, 32- and 64-wide heads and the hidden-384 (fused LayerNorm) path."""
    import os
    from qre_b200 import _cabi
    g = np.load(os.path.join(golden_dir, "embedder.npz"))
    for name in ("hd32", "hd64", "h384"):
        V, H, L, nh, F, P, S, seed = (int(x) for x in g[name + "_cfg"])
        o = MO.EncoderConfig(vocab_size=V, hidden=H, layers=L, heads=nh, intermediate=F, max_positions=P, max_seq_len=S)
        lens = g[name + "_lens"]
        off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
        enc = _cabi.Encoder(_cabi.EncoderConfig(V, H, L, nh, F, P, 2, 1e-12, S), MO.random_weights(o, seed=seed),
                            max_tokens=int(off[-1]) + 64, max_seqs=len(lens))
        got = enc.encode_packed(g[name + "_ids"], off)
        _assert_close(got, g[name + "_emb"], lens)
        raw = enc.encode_packed(g[name + "_ids"], off, normalize=False)
        assert np.abs(raw - g[name + "_pooled"]).max() <= 2e-2 * np.abs(g[name + "_pooled"]).max(), name
        enc.close()


def test_small_encoder_head_dim_32():
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=500, hidden=128, layers=2, heads=4, intermediate=256, max_positions=64,
                         max_seq_len=48)
    c = _cabi.EncoderConfig(500, 128, 2, 4, 256, 64, 2, 1e-12, 48)
    lens = [1, 2, 7, 15, 16, 17, 31, 32, 33, 48, 0, 60, 5]
    got, want = _check(o, c, lens, seed=3)
    _assert_close(got, want, lens)


def test_small_encoder_head_dim_64():
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=300, hidden=128, layers=3, heads=2, intermediate=192, max_positions=128,
                         max_seq_len=100)
    c = _cabi.EncoderConfig(300, 128, 3, 2, 192, 128, 2, 1e-12, 100)
    lens = [3, 100, 64, 65, 1, 18]
    got, want = _check(o, c, lens, seed=4)
    _assert_close(got, want, lens)


def test_bert_base_width_head_dim_64():
    """hidden 768 / 12 heads / FFN 3072 (all-mpnet-base / BERT-base widths), two layers."""
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=1000, hidden=768, layers=2, heads=12, intermediate=3072, max_positions=128,
                         max_seq_len=128)
    c = _cabi.EncoderConfig(1000, 768, 2, 12, 3072, 128, 2, 1e-12, 128)
    lens = [5, 64, 65, 128, 17, 1, 90]
    got, want = _check(o, c, lens, seed=8)
    _assert_close(got, want, lens)


def test_minilm_l6_shape_random_weights():
    """The all-MiniLM-L6-v2 architecture (qre:125) with seeded random weights, ragged lengths 1..256
    (max_seq_length) and one over-long sequence that must be truncated."""
    from qre_b200 import _cabi
    lens = [1, 2, 3, 8, 9, 16, 24, 31, 32, 33, 47, 64, 100, 128, 129, 200, 255, 256, 300, 12, 12, 0, 5]
    got, want = _check(MO.MINILM_L6, _cabi.EncoderConfig(), lens, seed=11)
    _assert_close(got, want, lens)


def test_unnormalised_pool_and_batch_independence():
    """normalize=0 returns the mean-pooled vector; a sequence's embedding does not depend on what else
    is in the batch (bit-identical alone and inside a batch)."""
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=500, hidden=128, layers=2, heads=4, intermediate=256, max_positions=64,
                         max_seq_len=48)
    c = _cabi.EncoderConfig(500, 128, 2, 4, 256, 64, 2, 1e-12, 48)
    lens = [9, 30, 4]
    got, want = _check(o, c, lens, seed=5, normalize=False)
    assert np.abs(got - want).max() <= 2e-2 * np.abs(want).max()
    w = MO.random_weights(o, seed=5)
    rs = np.random.Generator(np.random.PCG64(105))
    seqs = [rs.integers(0, 500, n).astype(np.int32) for n in lens]
    enc = _cabi.Encoder(c, w, max_tokens=256, max_seqs=8)
    whole = enc.encode_ids(seqs)
    for i, s in enumerate(seqs):
        assert np.array_equal(enc.encode_ids([s])[0], whole[i])
    # capacity and argument errors come back as errors, not crashes
    with pytest.raises(_cabi.QreError, match="capacity"):
        enc.encode_ids([np.zeros(40, np.int32)] * 9)
    with pytest.raises(_cabi.QreError, match="tokens"):
        enc.encode_ids([np.zeros(48, np.int32)] * 6)
    enc.close()


@pytest.mark.parametrize("H,heads,F,P,lens", [
    (1024, 16, 4096, 512, [512, 300, 1, 77]),       # the documented maxima: hidden, FFN, sequence length, 64-wide heads
    (64, 2, 64, 32, [1, 5, 32, 16, 0, 40]),         # the minima: one 192-feature tile mostly clipped
    (320, 10, 1280, 96, [96, 33, 64, 65, 7]),       # widths that are not multiples of the 192-feature tile
])
def test_extreme_shapes(H, heads, F, P, lens):
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=300, hidden=H, layers=1, heads=heads, intermediate=F, max_positions=P,
                         max_seq_len=P)
    c = _cabi.EncoderConfig(300, H, 1, heads, F, P, 2, 1e-12, P)
    got, want = _check(o, c, lens, seed=H)
    _assert_close(got, want, [min(n, P) for n in lens])


def test_small_and_large_batches_agree():
    """hidden 384: FFN-down runs with LayerNorm fused in its epilogue from ~9.5 k tokens up and with the
    separate LayerNorm kernel below.  The same sequences embedded alone (small) and inside a 12 k-token
    batch (large) must agree to bf16 noise, and the large batch must itself meet the oracle tolerance."""
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=800, hidden=384, layers=2, heads=12, intermediate=1536, max_positions=64,
                         max_seq_len=64)
    c = _cabi.EncoderConfig(800, 384, 2, 12, 1536, 64, 2, 1e-12, 64)
    w = MO.random_weights(o, seed=13)
    rs = np.random.Generator(np.random.PCG64(14))
    lens = rs.integers(8, 64, 340)
    seqs = [rs.integers(0, 800, n).astype(np.int32) for n in lens]
    assert sum(lens) > 10500
    enc = _cabi.Encoder(c, w, max_tokens=int(sum(lens)) + 64, max_seqs=len(seqs))
    large = enc.encode_ids(seqs)
    small = enc.encode_ids(seqs[:6])
    cos = (large[:6] * small).sum(1)
    assert cos.min() > 0.9999, float(cos.min())
    assert np.abs(large[:6] - small).max() < 4e-3
    want = MO.encode_ids(w, seqs[:24], o)
    _assert_close(large[:24], want, lens[:24])
    enc.close()


def test_encode_dev_feeds_the_lookup_without_a_host_round_trip():
    """ids -> qre_encode_dev -> qre_cache_append_dev / lookup on device pointers: a text inserted through
    the encoder is found again by the same text (sim ~ 1) and by nothing else."""
    import torch
    from qre_b200 import _cabi
    o = MO.EncoderConfig(vocab_size=500, hidden=128, layers=2, heads=4, intermediate=256, max_positions=64,
                         max_seq_len=48)
    c = _cabi.EncoderConfig(500, 128, 2, 4, 256, 64, 2, 1e-12, 48)
    w = MO.random_weights(o, seed=6)
    rs = np.random.Generator(np.random.PCG64(9))
    lens = rs.integers(3, 40, 64)
    seqs = [rs.integers(0, 500, n).astype(np.int32) for n in lens]
    off = np.zeros(65, np.int32)
    off[1:] = np.cumsum(lens)
    dev = torch.device("cuda", 0)
    ids_d = torch.from_numpy(np.concatenate(seqs)).to(dev)
    off_d = torch.from_numpy(off).to(dev)
    out_d = torch.empty((64, 128), dtype=torch.float32, device=dev)
    enc = _cabi.Encoder(c, w, max_tokens=4096, max_seqs=64)
    torch.cuda.synchronize()
    enc.encode_dev(64, ids_d.data_ptr(), off_d.data_ptr(), int(off[-1]), int(lens.max()), out_d.data_ptr(), stream=1)
    torch.cuda.synchronize()
    host = enc.encode_ids(seqs)
    assert np.array_equal(out_d.cpu().numpy(), host)
    from qre_b200.sharded import ShardedLookup
    sh = _cabi.Shard(dim=128, capacity=1024, max_batch=64)
    n = 64
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    ctx_d = t(np.full(n, 2, np.int64))
    ts_d, lru_d = t(np.zeros(n, np.int64)), t(np.arange(1, n + 1, dtype=np.int64))
    pd_d, pc_d = t(np.zeros(n, np.uint8)), t(np.ones(n))
    torch.cuda.synchronize()
    slots = sh.append_dev(n, out_d.data_ptr(), ctx_d.data_ptr(), ts_d.data_ptr(), lru_d.data_ptr(), pd_d.data_ptr(),
                          pc_d.data_ptr())
    assert np.array_equal(slots, np.arange(n))
    sh.set_centroids(host[:2].copy(), np.array([0, 1], np.uint8))
    torch.cuda.synchronize()
    look = ShardedLookup(sh, _cabi.RouteConfig(ttl_us=10**12))
    # a second encode of the same ids, straight into the lookup
    q_d = torch.empty_like(out_d)
    enc.encode_dev(64, ids_d.data_ptr(), off_d.data_ptr(), int(off[-1]), int(lens.max()), q_d.data_ptr(), stream=1)
    out = look.step(q_d, ctx_d, 1, t(np.full(n, 10, np.int32)), t(np.zeros(n, np.uint8)), t(np.full(n, 0.9)))
    torch.cuda.synchronize()
    assert np.array_equal(out["hit_slot"].cpu().numpy(), np.arange(n))
    assert np.all(out["hit_sim"].cpu().numpy() > 0.9999)
    sh.close()
    enc.close()
