/*
 * qre_encoder_b200.h — C ABI of the on-GPU query embedder in libqre_b200.so.
 *
 * What it replaces.  The reference embeds every query on the host right before the lookup:
 *     q_emb = self.cache_embedder.encode([query])[0]        src/query_router_engine.py:571
 *     q_emb = self.embedder.encode([query])[0]              src/query_router_engine.py:181
 *     vecs  = embedder.encode(texts)                        src/cache.py:412  (warm_up)
 * where the embedder is SentenceTransformer("all-MiniLM-L6-v2") (qre:125-126, 510-511):
 * WordPiece tokenisation -> BertModel (6 layers, hidden 384, 12 heads, FFN 1536) -> mean over
 * tokens -> L2 normalise.  The algorithm lives in sentence-transformers 3.4.1 / transformers
 * (requirements.txt:52-53), not in the reference's tree; oracle/minilm_oracle.py restates it
 * and is pinned against transformers' BertModel / BertTokenizer.
 *
 * Here the encoder runs on the same GPU as the cache: token ids go in, fp32 [m, hidden]
 * embeddings come out in device memory, laid out exactly as qre_scan_dev / qre_cache_append_dev
 * take them, so a batch of queries never returns to the host between "tokenised" and "routed".
 * GEMMs are tcgen05 bf16 -> fp32 kernels with the bias / GELU / residual epilogues fused;
 * attention runs per (sequence, head) on mma.sync tensor cores; LayerNorm, pooling and the
 * final normalisation are fp32.  Sequences are packed (no padding): `offsets[m+1]` delimits
 * each query's ids inside one flat array.
 *
 * One batch at a time per encoder handle: it owns one set of activation buffers, so calls from several
 * threads or on several streams must be ordered by the caller (the Python class holds a lock).
 *
 * Conventions are those of qre_b200.h: int status codes (0 = OK), qre_last_error() for the
 * message, no exceptions, caller owns host buffers, no CPU fallback.
 */
#ifndef QRE_ENCODER_B200_H
#define QRE_ENCODER_B200_H

#include "qre_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qre_encoder qre_encoder_t;
typedef struct qre_wordpiece qre_wordpiece_t;

/* Shape of the BERT-style encoder.  all-MiniLM-L6-v2: {30522, 384, 6, 12, 1536, 512, 2, 1e-12f, 256}.
 * Supported: hidden % 64 == 0 and <= 1024, intermediate % 64 == 0 and <= 4096,
 * hidden / heads in {32, 64}, max_seq_len <= min(512, max_positions). */
typedef struct {
  int32_t vocab_size;
  int32_t hidden;
  int32_t layers;
  int32_t heads;
  int32_t intermediate;
  int32_t max_positions;
  int32_t type_vocab;
  float ln_eps;
  int32_t max_seq_len;     /* longer sequences are truncated (sentence-transformers max_seq_length) */
  int32_t reserved;
} qre_encoder_config_t;

/* One encoder layer's parameters, host fp32, in torch.nn.Linear layout ([out, in] row-major).
 * Names follow transformers' BertModel.state_dict(): attention.self.{query,key,value},
 * attention.output.{dense,LayerNorm}, intermediate.dense, output.{dense,LayerNorm}. */
typedef struct {
  const float *wq, *bq, *wk, *bk, *wv, *bv;      /* [H,H], [H] each */
  const float *wo, *bo, *ln1_g, *ln1_b;          /* [H,H], [H], [H], [H] */
  const float *w1, *b1;                          /* [F,H], [F] */
  const float *w2, *b2, *ln2_g, *ln2_b;          /* [H,F], [H], [H], [H] */
} qre_encoder_layer_t;

/* Allocates weights and activation buffers for up to max_tokens packed tokens in up to
 * max_seqs (<= 65535) sequences per call on `device`.  Replaces SentenceTransformer(model_name)
 * (qre:126, 511) minus the checkpoint download: weights are supplied by the caller. */
QRE_API int qre_encoder_create(int device, const qre_encoder_config_t* cfg, int64_t max_tokens, int max_seqs,
                               qre_encoder_t** out);
QRE_API int qre_encoder_destroy(qre_encoder_t* e);
/* embeddings.{word,position,token_type}_embeddings.weight and embeddings.LayerNorm.{weight,bias} */
QRE_API int qre_encoder_set_embeddings(qre_encoder_t* e, const float* word, const float* position,
                                       const float* token_type, const float* ln_g, const float* ln_b);
QRE_API int qre_encoder_set_layer(qre_encoder_t* e, int layer, const qre_encoder_layer_t* w);

/* encode(): host ids in, host embeddings out.  ids = concatenation of the m sequences,
 * offsets[i]..offsets[i+1] delimits sequence i (offsets[0] == 0; int32).  Sequences longer than
 * max_seq_len are truncated; an empty sequence yields a zero row.  out: [m, hidden] fp32.
 * normalize != 0 applies the model's final Normalize module (what encode() returns for
 * all-MiniLM-L6-v2); 0 returns the mean-pooled vector. */
QRE_API int qre_encode(qre_encoder_t* e, int m, const int32_t* ids, const int32_t* offsets, int normalize,
                       float* out);
/* Device variant: ids / offsets / out are device pointers, nothing is copied or synchronised;
 * n_tokens = offsets[m] and max_len = the longest sequence (after truncation the caller applied:
 * every sequence must already satisfy len <= max_seq_len).  `out` can be handed straight to
 * qre_scan_dev / qre_scan_publish_dev / qre_cache_append_dev.  `stream`: cudaStream_t, NULL =
 * the encoder's own stream. */
QRE_API int qre_encode_dev(qre_encoder_t* e, int m, const int32_t* ids, const int32_t* offsets,
                           int64_t n_tokens, int max_len, int normalize, float* out, void* stream);
/* Kernel launches issued so far, and CUDA-event timing of the GEMM kernels (bench tooling; no
 * counterpart in the reference). */
QRE_API int qre_encoder_launch_count(const qre_encoder_t* e, int64_t* count);

/* ---- bring-up / test hooks (no counterpart in the reference) --------------------------------
 * C[rows, n] = epilogue(A[rows, k] * W[n, k]^T + bias) through the encoder's tcgen05 GEMM kernel,
 * host fp32 in/out (A and W are rounded to bf16 on the way in, C is bf16 rounded on the way
 * out).  epilogue: 0 = bias, 1 = bias + exact GELU, 2 = bias + residual[rows, n],
 * 3 = LayerNorm(bias + residual) with ln_g / ln_b / ln_eps (n must be 384; NULL / 0 otherwise). */
QRE_API int qre_encoder_debug_gemm(int device, int rows, int n, int k, const float* a, const float* w,
                                   const float* bias, const float* residual, int epilogue, const float* ln_g,
                                   const float* ln_b, float ln_eps, float* c);

/* ---- WordPiece on host threads --------------------------------------------------------------
 * BERT uncased tokenisation (BasicTokenizer + greedy longest-match-first WordPiece), the step
 * SentenceTransformer.encode runs in Python before the model.  `vocab_path`: a vocab.txt, one
 * piece per line.  ASCII texts take a byte-wise path; other texts are decoded, cleaned, NFD-
 * decomposed, stripped of accents and lower-cased through tables generated from Python's unicodedata
 * and cross-checked against the Rust `tokenizers` backend (csrc/qre_unicode_tables.inc).  Texts holding
 * a character whose handling depends on its neighbours (capital sigma), on which those two reference
 * implementations disagree (unassigned code points, a few hundred recent characters) or malformed
 * UTF-8 are NOT tokenised here: their length is reported as -1 and the caller routes them to its
 * own tokenizer. */
QRE_API int qre_wordpiece_create(const char* vocab_path, int do_lower_case, qre_wordpiece_t** out);
QRE_API int qre_wordpiece_destroy(qre_wordpiece_t* t);
/* ids_out: capacity `cap` ids in total; lens_out[n]: ids written for text i ([CLS] .. [SEP]
 * included, truncated to max_len), or -1 for a text handed back.  Sequence i starts at the sum of
 * the non-negative lens before it.  n_threads <= 0: all host cores. */
QRE_API int qre_wordpiece_encode(const qre_wordpiece_t* t, int64_t n, const char* const* texts,
                                 const int64_t* text_len, int max_len, int32_t* ids_out, int64_t cap,
                                 int32_t* lens_out, int n_threads);
/* The same with the texts in ONE byte blob: text i = blob[offsets[i] .. offsets[i+1]), n + 1 offsets
 * (what a Python caller builds with one join + encode instead of n byte strings and n pointers). */
QRE_API int qre_wordpiece_encode_packed(const qre_wordpiece_t* t, int64_t n, const char* blob,
                                        const int64_t* offsets, int max_len, int32_t* ids_out, int64_t cap,
                                        int32_t* lens_out, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* QRE_ENCODER_B200_H */
