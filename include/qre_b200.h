/* qre_b200.h — C ABI of libqre_b200.so: the B200 (sm_100a) routing-cache hot path.
 *
 * The reference (clumpygum/distributed-llm) has no FFI: its boundary for this path is the
 * Python class API (`QueryCache` in src/cache.py, `SemanticRouter` / `HybridRouter` /
 * `QueryRouter` in src/query_router_engine.py).  Each entry point below names the reference
 * code it replaces (file:line relative to /root/reference/).  The Python mirror in
 * `distributed-llm_b200/` binds these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions: plain pointers and sizes only; every function returns an int status
 * (QRE_OK = 0); no exceptions cross the ABI; `qre_last_error()` gives the message of the
 * calling thread's last failure.  Functions without the `_dev` suffix take HOST buffers and
 * are synchronous (copies in, kernels, copies out).  `_dev` functions take DEVICE pointers
 * (on the handle's GPU), enqueue on `stream` (a cudaStream_t; NULL = the handle's own stream,
 * so pass cudaStreamLegacy to name the default stream) and return without synchronising.  A handle is not thread-safe: serialise calls per handle (the
 * reference holds one RLock around its store, src/cache.py:225).
 *
 * There is no CPU fallback: `qre_cache_create` fails with QRE_ERR_NO_DEVICE when no sm_100
 * GPU is present.
 */
#ifndef QRE_B200_H
#define QRE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QRE_ABI_VERSION 2

#if defined(__GNUC__)
#define QRE_API __attribute__((visibility("default")))
#else
#define QRE_API
#endif

enum {
  QRE_OK = 0,
  QRE_ERR_INVALID = 1,     /* bad argument */
  QRE_ERR_CUDA = 2,        /* a CUDA call failed; see qre_last_error() */
  QRE_ERR_CAPACITY = 3,    /* cache full / batch larger than max_batch */
  QRE_ERR_NO_DEVICE = 4,   /* no usable sm_100 device */
  QRE_ERR_OVERFLOW = 5     /* candidate window overflow could not be resolved (no longer returned: every
                              overflowed query is settled on the device; kept for ABI stability) */
};

#define QRE_TOPK 8                 /* candidates kept per query (per shard) */
#define QRE_MAX_DEVICES 32         /* D: routing targets */
#define QRE_MAX_CENTROIDS 256      /* K: topic centroids */
#define QRE_CTX_TOMBSTONE 0xFFFFFFFFFFFFFFFFull

/* scan_mode */
#define QRE_SCAN_TENSOR 0          /* tcgen05 bf16 GEMM + fused top-k, exact fp32 re-score */
#define QRE_SCAN_EXACT  1          /* brute-force fp32 scan (fallback / cross-check) */

/* semantic method codes (src/query_router_engine.py:190,201,210) */
#define QRE_SEM_SEMANTIC   0
#define QRE_SEM_IRRELEVANT 1
#define QRE_SEM_AMBIGUOUS  2

/* result flags (per query) */
#define QRE_FLAG_ZERO_NORM   1     /* ||q|| < 1e-9: lookup returns miss (src/cache.py:284) */
#define QRE_FLAG_OVERFLOW    2     /* the batch's spill pool ran out while this query still had contenders;
                                      settled by the brute-force fp32 pass: the winner is exact, but cand[]
                                      then lists the winner only */

typedef struct qre_cache qre_cache_t;

/* One exactly re-scored candidate; also the unit of the multi-GPU all-gather (32 bytes). */
typedef struct {
  float    sim;        /* fl32(sdot32(q,e) / fl32(|q|*|e|))  — src/cache.py:291 */
  uint32_t shard;      /* rank that owns the row */
  int64_t  slot;       /* row slot within that shard; -1 = empty */
  uint64_t lru_seq;    /* OrderedDict position: smaller = earlier (tie-break, src/cache.py:292) */
  double   pred_conf;  /* stored predict_device() confidence — src/cache.py:106-140 */
} qre_candidate_t;
/* pred_dev travels in the top byte of `shard` (shard = rank | pred_dev << 24). */

/* Parameters of one lookup+route pass.  Defaults in comments = the shipped configs
 * (src/query_router_engine.py:704-731). */
typedef struct {
  double  similarity_threshold;     /* 0.85  — src/cache.py:214, compared in double (cache.py:292) */
  int64_t now_us;                   /* one `now` per batch, microseconds (src/cache.py:524) */
  int64_t ttl_us;                   /* ttl_seconds * 1e6 */
  int32_t scan_mode;                /* QRE_SCAN_* */
  int32_t do_route;                 /* 0 = lookup only (QueryCache.lookup), 1 = also route */
  int32_t n_devices;                /* D; reference = 2 (0 = nano, 1 = orin) */
  int32_t token_threshold;          /* 1000 — qre:87 */
  double  w_token, w_semantic, w_heuristic;  /* 0.25 / 0.45 / 0.30 — qre:539,715 */
  double  w_perf;                   /* 0 in the reference; added term w_perf * H[dev] */
  double  sem_min_similarity;       /* 0.05 — qre:129 */
  double  sem_margin_threshold;     /* 0.03 — qre:128 */
  int32_t pipeline_slot;            /* 0 or 1: which copy of the per-batch query state this call uses.
                                       qre_scan_dev(slot) and the qre_merge_route_dev(slot) of the SAME batch
                                       pass the same value; alternate it between batches to overlap the merge of
                                       batch i (another stream) with the scan of batch i+1.  0 otherwise. */
  int32_t reserve_sms;              /* SMs the scan leaves free for kernels running beside it (0 = none) */
  int32_t slice_results;            /* Row-sharded cache (qre_xchg_* connected) only, through qre_submit_dev and
                                       qre_lookup_route_begin / _end.  0: every rank merges and routes all m
                                       queries (replicated results).  1: rank r merges and routes only queries
                                       [r*per, min(m, (r+1)*per)), per = ceil(m / world) — the candidate
                                       exchange becomes an all-to-all of slices, the output arrays stay m long
                                       and only that slice of them is written, and the host-buffer call uploads
                                       only that slice of the query matrix (the ranks exchange the slices over
                                       NVLink).  Ignored without an exchange. */
  int32_t reserved;                 /* 0 */
} qre_params_t;

/* Per-query outputs, struct-of-arrays.  Any pointer may be NULL (not wanted).
 * Host pointers for qre_lookup_route, device pointers for the _dev functions. */
typedef struct {
  int64_t* hit_slot;        /* [m]      -1 = miss           — src/cache.py:293-296 */
  float*   hit_sim;         /* [m]      0 on a miss */
  uint8_t* hit_pred_dev;    /* [m]      stored prediction of the hit row — cache.py:311 */
  double*  hit_pred_conf;   /* [m] */
  uint32_t* hit_shard;      /* [m]      owning rank of the hit row */
  qre_candidate_t* cand;    /* [m, QRE_TOPK] merged candidates, best first */
  float*   centroid_sim;    /* [m, K]   qre:133-138 */
  uint8_t* tok_dev;         /* [m]      qre:98 */
  double*  tok_conf;        /* [m]      qre:99 */
  uint8_t* sem_dev;         /* [m]      qre:180-213 */
  double*  sem_conf;        /* [m] */
  uint8_t* sem_method;      /* [m]      QRE_SEM_* */
  uint8_t* dev;             /* [m]      hybrid winner — qre:401-404 */
  double*  conf;            /* [m]      qre:406-407 */
  double*  score;           /* [m, D]   S[dev] — qre:389-398 */
  uint8_t* flags;           /* [m]      QRE_FLAG_* */
  /* heur_dev[i] must be < n_devices (larger values are clamped to the last device); with
   * n_devices == 1 the token router's "large" answer (device 1, qre:98) is clamped to device 0. */
} qre_result_t;

/* ---- library ------------------------------------------------------------------------- */
QRE_API int qre_abi_version(void);
QRE_API const char* qre_last_error(void);
QRE_API int qre_device_count(int* count);

/* ---- cache object: replaces QueryCache's store (src/cache.py:224) ---------------------
 * dim: embedding dimension (multiple of 8); capacity: max rows on this GPU (QueryCache
 * max_size / number of shards, src/cache.py:212); max_batch: largest m a lookup may pass. */
QRE_API int qre_cache_create(int device, int dim, int64_t capacity, int max_batch, qre_cache_t** out);
QRE_API int qre_cache_destroy(qre_cache_t* h);
QRE_API int qre_cache_info(const qre_cache_t* h, int64_t* capacity, int64_t* high_water, int64_t* live,
                   int* dim, int* max_batch, int* device);
/* The handle's CUDA stream (a cudaStream_t) for interop. */
QRE_API int qre_cache_stream(const qre_cache_t* h, void** stream);

/* Insert n new rows (QueryCache.insert new-hash branch, src/cache.py:356-370): copies the
 * fp32 rows, stores |e|, the L2-normalised bf16 copy and the row metadata; returns the
 * slots chosen (tombstoned slots are reused first). */
QRE_API int qre_cache_append(qre_cache_t* h, int64_t n, const float* rows, const uint64_t* ctx,
                     const int64_t* ts_us, const uint64_t* lru_seq, const uint8_t* pred_dev,
                     const double* pred_conf, int64_t* slot_out);
QRE_API int qre_cache_append_dev(qre_cache_t* h, int64_t n, const float* rows, const uint64_t* ctx,
                         const int64_t* ts_us, const uint64_t* lru_seq, const uint8_t* pred_dev,
                         const double* pred_conf, int64_t* slot_out_host, void* stream);
/* Refresh-in-place (src/cache.py:345-354) and LRU promotion (cache.py:261,303): overwrite
 * row data and/or metadata of existing slots.  NULL arrays are left unchanged. */
QRE_API int qre_cache_update(qre_cache_t* h, int64_t n, const int64_t* slot, const float* rows,
                     const uint64_t* ctx, const int64_t* ts_us, const uint64_t* lru_seq,
                     const uint8_t* pred_dev, const double* pred_conf);
/* Delete rows (src/cache.py:526-530 _delete_entry; eviction cache.py:532-554). */
QRE_API int qre_cache_tombstone(qre_cache_t* h, int64_t n, const int64_t* slot);
QRE_API int qre_cache_clear(qre_cache_t* h);                          /* src/cache.py:502-512 */
/* Read back fp32 rows (for QueryCache.save, src/cache.py:426-432). */
QRE_API int qre_cache_read_rows(qre_cache_t* h, int64_t n, const int64_t* slot, float* rows_out);

/* Binary snapshot of the device-resident rows and row metadata, beside the JSON save/load the
 * reference offers (src/cache.py:426-465; 1M x 384 floats are ~8 GB of JSON text).  Load needs an
 * empty cache of the same dim and enough capacity; slots keep their numbers. */
QRE_API int qre_cache_snapshot_save(qre_cache_t* h, const char* path);
QRE_API int qre_cache_snapshot_load(qre_cache_t* h, const char* path);

/* ---- router state: replaces SemanticRouter centroids (qre:131,140-178) ---------------- */
QRE_API int qre_set_centroids(qre_cache_t* h, int n_centroids, const float* centroids,
                      const uint8_t* owner_device);
QRE_API int qre_set_perf_health(qre_cache_t* h, int n_devices, const double* health);

/* ---- the hot path ---------------------------------------------------------------------
 * QueryCache.lookup semantic scan (src/cache.py:267-305) for m queries against one snapshot
 * of the cache, fused with SemanticRouter.route (qre:180-213), TokenBasedRouter.route
 * (qre:98-99) and HybridRouter.route (qre:382-414).  `tok`, `heur_dev`, `heur_conf` are the
 * host-computed per-query features (token count; HeuristicRouter result, qre:323-364); they
 * may be NULL when params->do_route == 0.  shard/n_shards describe this handle's place in a
 * row-sharded cache (0/1 for a single GPU). */
QRE_API int qre_lookup_route(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx,
                     const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                     const qre_params_t* params, qre_result_t* out);

/* The same call split in two so that consecutive batches overlap: `begin` enqueues the H2D
 * copy of the inputs (own copy stream), the kernels and the D2H copy of the results (own copy
 * stream) and returns; `end` waits for that batch and fills `out`.  Three slots (0 / 1 / 2): begin
 * batch i+1 on another slot before ending batch i and its input copy runs under batch i's
 * scan; with several ranks of a row-sharded cache (qre_xchg_*) keep two batches ahead (begin i+2
 * before end i), so that one rank's late host thread does not stall the scans of all of them.
 * The inputs must stay unchanged until `end`; if `queries` is page-locked memory it is copied
 * straight from the caller's buffer.  qre_lookup_route == begin(0) + end(0). */
QRE_API int qre_lookup_route_begin(qre_cache_t* h, int slot, int m, const float* queries,
                                   const uint64_t* q_ctx, const int32_t* tok, const uint8_t* heur_dev,
                                   const double* heur_conf, const qre_params_t* params);
QRE_API int qre_lookup_route_end(qre_cache_t* h, int slot, qre_result_t* out);

/* The whole pass on DEVICE pointers, pipelined inside the library (no reference counterpart: the
 * reference handles one query at a time on the host; this is the call a device-resident producer,
 * e.g. the on-GPU embedder, feeds).  qre_submit_dev enqueues one batch and returns: query
 * preparation runs on an auxiliary stream of the handle, the tensor-core scans ALONE on the handle's
 * main stream and a twin of it in turn (consecutive scans overlap only where one retires its CTAs
 * and the next moves in), and exact re-score, overflow settlement, the candidate exchange (when qre_xchg_* is
 * connected: every rank then submits the same batch) and merge + routing on a third stream, so
 * consecutive scans run back to back while the small kernels of the neighbouring batches run
 * beside them (leave them room with params->reserve_sms).  Slots 0 / 1 / 2 name the caller's output
 * arrays: submit batch i on slot i & 1 (or i % 3); a slot's previous batch is waited for on the
 * device, not on the host.  Cache rows must not be written (append / update / tombstone / clear)
 * while a submitted batch is in flight: wait for it first.  The inputs must be valid
 * on `in_stream` at the time of the call (NULL: already valid) and, like *out_dev's arrays, stay
 * untouched until the batch has finished.  qre_wait_dev makes `stream` wait for the batch last
 * submitted on `slot` (stream == NULL: blocks the host instead). */
QRE_API int qre_submit_dev(qre_cache_t* h, int slot, int m, const float* queries, const uint64_t* q_ctx,
                           const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                           const qre_params_t* params, qre_result_t* out_dev, void* in_stream);
QRE_API int qre_wait_dev(qre_cache_t* h, int slot, void* stream);

/* One process driving several GPUs: the same call over a GROUP of handles, one per GPU, that
 * together hold one row-sharded cache (the single-process shape of SURVEY §8b/§8e; the reference
 * builds its one cache at src/query_router_engine.py:497-505).  Every handle scans and exactly
 * re-scores its own rows for the whole batch on its GPU, the per-shard candidate lists are copied
 * to the first handle's GPU and merged there with the reference's comparator (sim desc, LRU order
 * asc, src/cache.py:292), followed by the routing epilogue.  Host buffers, synchronous.
 * out->hit_shard[i] = index in hs[] of the handle that holds the hit row, out->hit_slot[i] its slot
 * there.  Centroids / perf health are read from hs[0]. */
QRE_API int qre_group_lookup_route(qre_cache_t* const* hs, int n_handles, int m, const float* queries,
                                   const uint64_t* q_ctx, const int32_t* tok, const uint8_t* heur_dev,
                                   const double* heur_conf, const qre_params_t* params, qre_result_t* out);

/* Device-pointer pipeline, one stage per call (all enqueue on `stream`):
 *   qre_scan_dev   : queries -> this shard's exactly re-scored candidates cand[m, QRE_TOPK]
 *   (multi-GPU: all-gather cand over NCCL, e.g. torch.distributed.all_gather_into_tensor)
 *   qre_merge_route_dev : n_shards candidate lists -> hit decision + routing outputs.
 * qre_scan_dev also writes *overflow_count (device int32): number of queries whose candidate
 * window overflowed and were NOT resolved on the device; if it is non-zero the caller must
 * call qre_resolve_overflow_dev before merging. */
QRE_API int qre_scan_dev(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx,
                 const qre_params_t* params, int shard, qre_candidate_t* cand_out,
                 int32_t* overflow_count, void* stream);
QRE_API int qre_resolve_overflow_dev(qre_cache_t* h, int m, const qre_params_t* params, int shard,
                             qre_candidate_t* cand_inout, void* stream);
QRE_API int qre_merge_route_dev(qre_cache_t* h, int m, int n_shards, const qre_candidate_t* cand_all,
                        const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                        const qre_params_t* params, qre_result_t* out_dev, void* stream);

/* ---- host-side per-query features for a batch (no GPU involved) -----------------------------
 * HeuristicRouter.route (qre:323-364, patterns qre:231-294) and the no-tokenizer token estimate
 * of TokenBasedRouter (qre:91,96) for n queries on `n_threads` C++ threads.  `queries[i]` are
 * UTF-8 bytes of length query_len[i]; context_len[i] is len(context) in characters (<= 0: no
 * context).  Outputs: tok[i]; heur_dev/heur_conf[i]; heur_rule[i] = which rule fired
 * (0 complex pattern, 1 long query, 2 multi-question, 3 code markers, 4 large context,
 * 5 simple pattern, 6 short everyday query, 7 token fallback; -1 = handed back to the caller's
 * own classifier: malformed UTF-8 only; every other text, ASCII or not, is classified with Python `re` /
 * `str` semantics for \b, \w, \s, \d, lower() incl. its final-sigma rule, strip(), split() and len()); heur_cat[i] = pattern category index for rules 0 / 5
 * (the '?' count for rule 2).  query_len / context_len: query_len[i] is in BYTES, context_len[i] in
 * characters. */
typedef struct {
  int32_t long_chars;            /* heuristic_long_chars          (qre:225) */
  int32_t multi_qmarks;          /* heuristic_multi_qmarks        (qre:226) */
  int32_t code_markers_needed;   /* heuristic_code_markers_needed (qre:227) */
  int32_t context_chars;         /* heuristic_context_chars       (qre:228) */
  int32_t token_threshold;       /* token_threshold               (qre:87)  */
  int32_t n_devices;
} qre_heur_params_t;
QRE_API int qre_host_features(int64_t n, const char* const* queries, const int64_t* query_len,
                              const int64_t* context_len, const qre_heur_params_t* hp, int n_threads,
                              int32_t* tok, uint8_t* heur_dev, double* heur_conf, int32_t* heur_rule,
                              int32_t* heur_cat);
/* The same with the queries in ONE byte blob (query i = blob[offsets[i] .. offsets[i+1]), n + 1
 * offsets), which spares a Python caller an array of n pointers; plus, when key_hex != NULL, the
 * cache's exact-match key of every ASCII query (QueryCache._make_hash, cache.py:518-521):
 * md5(ctx_keys[ctx_index[i]] + "||" + query.lower().strip()) as 32 lower-case hex characters at
 * key_hex + 32*i (zero bytes where heur_rule[i] == -1: the caller hashes those itself).  The n_ctx
 * distinct context keys are UTF-8 in ctx_blob with n_ctx + 1 offsets. */
QRE_API int qre_host_features_packed(int64_t n, const char* blob, const int64_t* offsets,
                                     const int64_t* context_len, const qre_heur_params_t* hp, int n_threads,
                                     int32_t* tok, uint8_t* heur_dev, double* heur_conf, int32_t* heur_rule,
                                     int32_t* heur_cat, const char* ctx_blob, const int64_t* ctx_offsets,
                                     const int32_t* ctx_index, int32_t n_ctx, char* key_hex);

/* ---- peer exchange: the all-gather written as stores into peer memory ------------------------
 * No counterpart in the reference (it never shards its cache; SURVEY §8e defines this exchange).
 * For a row-sharded cache with one process per GPU on one NVLink/NVSwitch box.  Instead of
 * handing cand[m, QRE_TOPK] to an NCCL all-gather, each rank's publish kernel writes its lists
 * directly into every rank's exchange buffer (CUDA-IPC-mapped peer pointers) and then raises a
 * per-rank flag there; the merge kernel of each rank waits for all flags of the batch before it
 * reads its local buffer.  Set-up: every rank calls qre_xchg_init (allocates its buffer, returns
 * a 64-byte IPC handle), the ranks exchange handles out of band (e.g. all_gather_object) and call
 * qre_xchg_connect once per peer.  Per batch: qre_scan_publish_dev then
 * qre_merge_route_xchg_dev, in that order and on ONE stream, the same number of times on every
 * rank: publish(e) must be followed by merge(e) before publish(e+1) (a second publish without the
 * merge in between is rejected — a peer two batches ahead would otherwise overwrite lists a merge
 * is still reading).  A peer that never publishes
 * makes the merge kernel trap after a few seconds instead of hanging.  Before destroying a
 * handle, synchronise all ranks (peers write into its buffer). */
#define QRE_IPC_HANDLE_BYTES 64
QRE_API int qre_xchg_init(qre_cache_t* h, int world, int rank, uint8_t* handle_out);
QRE_API int qre_xchg_connect(qre_cache_t* h, int peer, const uint8_t* handle);
/* qre_publish_dev alone publishes lists a previous qre_scan_dev left in `cand` (DEVICE pointer), so
 * the exchange can sit on another stream than the scan; qre_scan_publish_dev = scan + publish. */
QRE_API int qre_publish_dev(qre_cache_t* h, int m, const qre_candidate_t* cand, void* stream);
QRE_API int qre_scan_publish_dev(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx,
                                 const qre_params_t* params, int32_t* overflow_count, void* stream);
QRE_API int qre_merge_route_xchg_dev(qre_cache_t* h, int m, const int32_t* tok, const uint8_t* heur_dev,
                                     const double* heur_conf, const qre_params_t* params,
                                     qre_result_t* out_dev, void* stream);

/* Kernel launches issued by this handle since creation (bench.py's gpu_launches).  Measurement
 * tooling: no counterpart in the reference. */
QRE_API int qre_launch_count(const qre_cache_t* h, int64_t* count);

/* Measurement tooling, no counterpart in the reference.
 * Bracket every tensor-core scan launch with CUDA events on its own stream (enable != 0
 * resets the counters); qre_profile_read waits for the recorded launches and returns their
 * summed device time.  bench.py uses this for the live roofline of the dominant kernel. */
QRE_API int qre_profile(qre_cache_t* h, int enable);
QRE_API int qre_profile_read(qre_cache_t* h, double* scan_ms_total, int* n_launches);

/* ---- lookup without any allocation --------------------------------------------------------
 * No counterpart in the reference (its cache is a Python dict; it never manages device memory).
 *
 * For callers that are not allowed to allocate GPU memory at lookup time: a process that gave
 * all of HBM to its own pool at start-up, a request thread that cannot take the device-wide
 * stall cudaMalloc/cudaFree cause, a plugin whose host forbids allocations outside the host's
 * allocator.  This is the brute-force exact lookup (same masks, same fp32 rounding sequence
 * and therefore the same hit row and score as QRE_SCAN_EXACT) over arrays the CALLER owns,
 * all DEVICE pointers on `device`:
 *   rows[n_rows, dim] fp32 (16-byte aligned, dim % 4 == 0), row_ctx / row_ts_us / row_lru [n_rows]
 *   (ctx == QRE_CTX_TOMBSTONE marks a deleted row), queries[m, dim], q_ctx[m];
 *   outputs hit_row[m] (-1 = miss) and hit_sim[m] (0 on a miss).
 * It calls no cudaMalloc / cudaFree / allocating library routine.  Its only scratch (per-CTA
 * partial winners and a completion ticket) is a 197,120-byte static __device__ array inside
 * the library, which therefore occupies that much memory on every GPU that loads the library
 * until the process exits.  Limits: m <= QRE_NOALLOC_MAX_QUERIES, dim <= QRE_NOALLOC_MAX_DIM,
 * n_rows <= QRE_NOALLOC_MAX_ROWS; anything larger is rejected with QRE_ERR_CAPACITY and a
 * message in `err` (caller's buffer of err_len bytes, may be NULL); a `rows` pointer that is not
 * 16-byte aligned is rejected with QRE_ERR_INVALID before anything is launched.
 * All calls on a device share the array and it keeps the previous call's contents (the part
 * that must be zero is cleared at the start of each call), so calls that overlap on the same
 * GPU race with each other: issue them one after another.  The first call on each device
 * resolves the array's device address (cudaGetSymbolAddress); make that first call before
 * starting any stream capture.  Asynchronous: work is enqueued on `stream` (a cudaStream_t;
 * NULL = the default stream) and the outputs are valid once the stream has caught up. */
#define QRE_NOALLOC_MAX_QUERIES 8
#define QRE_NOALLOC_MAX_DIM 1024
#define QRE_NOALLOC_MAX_ROWS 2147483647ll
QRE_API int qre_exact_lookup_noalloc(int device, int64_t n_rows, int dim, const float* rows,
                                     const uint64_t* row_ctx, const int64_t* row_ts_us,
                                     const uint64_t* row_lru, int m, const float* queries,
                                     const uint64_t* q_ctx, double similarity_threshold, int64_t now_us,
                                     int64_t ttl_us, int64_t* hit_row, float* hit_sim, void* stream,
                                     char* err, size_t err_len);

/* ---- bring-up / test hooks (no counterpart in the reference) --------------------------------
 * Raw bf16 tensor-core product of the resident queries (from the last scan) with cache rows
 * [row0, row0+n): out[m, n] fp32, host buffer.  Used by tests to check the tcgen05/TMA
 * descriptors independently of the top-k logic. */
QRE_API int qre_debug_gemm(qre_cache_t* h, int m, int64_t row0, int n, float* out);

#ifdef __cplusplus
}
#endif
#endif /* QRE_B200_H */
