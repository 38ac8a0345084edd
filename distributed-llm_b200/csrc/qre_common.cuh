// Shared definitions of libqre_b200: device data layout, kernel argument blocks, launch
// wrappers.  See DESIGN.md for the layout rationale.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/qre_b200.h"

namespace qre {

constexpr int kBM = 128;          // queries per tile (UMMA M, one TMEM lane per query)
constexpr int kBN = 256;          // cache rows per tile (UMMA N)
constexpr int kBK = 64;           // bf16 elements per k-block (= one 128-byte swizzle row)
constexpr int kStages = 4;        // B-tile pipeline depth (4 x 32 KB)
constexpr int kMaxKB = 6;         // resident-A variant: dim_pad <= 384
constexpr int kMaxKBStream = 16;  // streaming-A variant: dim_pad <= 1024
constexpr int kTopK = QRE_TOPK;
constexpr uint32_t kNoRow = 0xFFFFFFFFu;
constexpr int kExactGroup = 8;        // queries per brute-force pass
constexpr int kPipeSlots = 3;         // copies of the per-batch state: the pipeline keeps up to three batches
                                      // in flight (prep of i+1, scan of i, re-score / merge of i-1) and a
                                      // slot is reused only three batches later
constexpr int kApiSlots = 3;          // host-buffer / output slots a caller may have in flight (begin/end, submit)
constexpr uint32_t kNoChunk = 0xFFFFFFFFu;    // empty spill chain
constexpr uint32_t kChunkLost = 0xFFFFFFFEu;  // spill pool exhausted: this (query, range) lost candidates
constexpr int kSpillPer = 15;         // candidates per spill chunk (128-byte chunks)
constexpr float kNormEps = 1e-9f;                 // src/cache.py:284,289 ; qre:136

// |approx - cosine| bound of the bf16 tensor-core score: two bf16 roundings (2^-9 each)
// under Cauchy-Schwarz = 2^-8, plus fp32 accumulation / normalisation slack.
constexpr float kApproxEps = 0.00390625f + 0.000244140625f;

struct __align__(16) RowMeta {    // per cache row, read only on the (rare) candidate path
  uint64_t ctx;                   // context id; QRE_CTX_TOMBSTONE = deleted
  int64_t ts_us;                  // insert/refresh time, microseconds
};

struct __align__(8) PartialEntry {  // approximate candidate kept by the fused top-k
  float v;                          // bf16 tensor-core cosine
  uint32_t row;                     // slot, kNoRow = empty
};

// Candidates that fall off a (query, row range) top-k list while still inside the exactness
// window are chained here instead of being dropped: each epilogue thread owns
// its chain, chunks come from a per-batch pool by one atomic bump per 15 candidates.
struct __align__(16) SpillChunk {
  uint32_t next;                    // older chunk of the same chain, kNoChunk = end
  uint32_t n;                       // entries used
  PartialEntry e[kSpillPer];
};
static_assert(sizeof(SpillChunk) == 128, "one chunk = one 128-byte line");

struct ExactBest {                  // per-(CTA, query) winner of the brute-force scan
  float sim;
  uint32_t row;
  uint64_t lru;
};

// Arguments of the tensor-core scan kernel.
struct GemmArgs {
  int m;                    // queries in the batch
  int mblocks;              // ceil(m / 128)
  int tiles_n;              // ceil(n_rows / 256)
  int splits;               // S: row ranges per M-block; work units = S * mblocks
  int num_kb;               // dim_pad / 64
  long long n_rows;         // high-water mark of the shard
  float floor_v;            // max(threshold, 0) - eps: nothing below can be a hit
  long long now_us, ttl_us;
  const uint64_t* q_ctx;    // [m]
  const RowMeta* meta;      // [capacity]
  PartialEntry* partial;    // [mblocks*128][S][kTopK]   (scratch)
  uint32_t* spill_head;     // [mblocks*128][S]  chain head per (query, range): kNoChunk / kChunkLost / index
  SpillChunk* spill_pool;   // [spill_cap]
  uint32_t* spill_next;     // bump counter, zeroed per batch
  uint32_t spill_cap;
  uint32_t* work_ctr;       // [2] CTA-pair scan: next work unit to claim, pairs finished (both 0 between launches)
  float* dump;              // debug: raw accumulators out[q * dump_ld + (row - dump_row0)]
  long long dump_ld, dump_row0;
};

struct ScanGeometry {
  int mblocks, tiles_n, splits, grid;
  bool pair = false;        // CTA-pair (cta_group::2) kernel chosen
};
ScanGeometry plan_scan(int m, long long n_rows, int num_sms);
cudaError_t gemm_topk_init();
cudaError_t gemm_topk_pair_init();
ScanGeometry plan_scan_pair(int m, long long n_rows, int num_sms);
struct DeviceState;
int scan_sms(const DeviceState& s, const qre_params_t& p);
// The geometry (and kernel variant) one scan launch will use; rescore must ask the same question.
ScanGeometry plan_for(const DeviceState& s, int m, long long n_rows, const qre_params_t& p);

// All device state of one cache shard.
struct DeviceState {
  int device = 0, num_sms = 0;
  int dim = 0, dim_pad = 0, max_batch = 0, mb_pad = 0;
  long long capacity = 0, cap_pad = 0;
  cudaStream_t stream = nullptr;
  // cache rows
  float* rows_f32 = nullptr;          // [capacity, dim_pad]  fp32 master (zero padded)
  __nv_bfloat16* rows_bf16 = nullptr; // [cap_pad, dim_pad]   L2-normalised, bf16
  float* row_norm = nullptr;          // [capacity]           |e| in fp32
  RowMeta* meta = nullptr;            // [capacity]
  uint64_t* lru = nullptr;            // [capacity]
  uint8_t* pred_dev = nullptr;        // [capacity]
  double* pred_conf = nullptr;        // [capacity]
  // Per-batch state.  Everything one in-flight batch reads or writes exists kPipeSlots times
  // ("pipeline slots"): query preparation of batch i+1 and re-score / merge / routing of batch
  // i-1 run on other streams beside the scan of batch i.  (The stage-per-call API addresses
  // slots 0 / 1 through qre_params_t.pipeline_slot.)
  struct Batch {
    float* q_f32 = nullptr;             // [mb_pad, dim_pad]  zero padded fp32 queries
    __nv_bfloat16* q_bf16 = nullptr;    // [mb_pad, dim_pad]  L2-normalised bf16 operand of the scan
    float* q_norm = nullptr;            // [mb_pad]
    uint64_t* q_ctx = nullptr;          // [max_batch] private copy of the batch's contexts
    uint8_t* q_flags = nullptr;         // [max_batch] QRE_FLAG_*
    PartialEntry* partial = nullptr;    // [partial_entries]  per (query, range) approximate top-k
    uint32_t* spill_head = nullptr;     // [partial_entries / kTopK]
    SpillChunk* spill_pool = nullptr;   // [spill_cap]
    uint32_t* counters = nullptr;       // [8]: 0 spill bump counter, 1 overflowed queries, 2 unresolved (zeroed per
                                        // batch by the prep kernel); 4, 5 the pair scan's work-unit counters (self-resetting)
    int32_t* ovf_list = nullptr;        // [max_batch]
    qre_candidate_t* cand = nullptr;    // [max_batch, kTopK] this shard's exactly re-scored candidates
    CUtensorMap tmap_q;                 // 128-row boxes over q_bf16
    int32_t* ovf_count() const { return reinterpret_cast<int32_t*>(counters + 1); }
    int32_t* ovf_unresolved() const { return reinterpret_cast<int32_t*>(counters + 2); }
  };
  Batch batch[kPipeSlots];
  int slot = 0;
  Batch& cur() { return batch[slot]; }
  const Batch& cur() const { return batch[slot]; }
  // scratch shared by both slots (users are serialised on one stream)
  long long partial_entries = 0;
  uint32_t spill_cap = 0;
  ExactBest* exact_partial = nullptr; // [kExactGroup * exact_grid]
  int exact_grid = 0;
  int settle_groups = 0;              // ceil(max_batch / kExactGroup): every query of a batch can be settled
  unsigned* settle_tickets = nullptr; // [settle_groups]
  ExactBest* settle_partial = nullptr;  // [settle_groups][kExactGroup][exact_grid]
  // router state
  float* centroids = nullptr;         // [K, dim_pad]
  float* cent_norm = nullptr;         // [K]
  uint8_t* cent_owner = nullptr;      // [K]
  int n_centroids = 0;
  double* perf_health = nullptr;      // [QRE_MAX_DEVICES]
  CUtensorMap tmap_e, tmap_eh;        // cache rows (256-row box / 128-row half box)
  bool allow_pair = true;               // QRE_SCAN_SINGLE_CTA=1 forces the single-CTA kernel
  // Small batches run prep -> scan -> re-score -> settle -> merge on ONE stream; with this set the
  // links after prep are launched with programmatic stream serialization (qre_encoder.cuh): each
  // kernel's CTAs are scheduled and set up while its predecessor drains, and block in pdl_wait().
  bool pdl = false;
  // optional CUDA-event bracket around every scan-kernel launch (bench.py's live roofline)
  bool prof_on = false;
  cudaEvent_t* prof_ev = nullptr;     // [2 * kProfMax]
  int prof_used = 0;
};
constexpr int kProfMax = 2048;


// ---- launch wrappers (qre_kernels.cu / qre_gemm_topk.cu); each returns #launches --------
int launch_append(DeviceState& s, long long n, const float* rows_dev, long long rows_ld, const int64_t* slot_dev,
                  const uint64_t* ctx, const int64_t* ts, const uint64_t* lru, const uint8_t* pdev,
                  const double* pconf, cudaStream_t st);
int launch_update_meta(DeviceState& s, long long n, const int64_t* slot_dev, const uint64_t* ctx,
                       const int64_t* ts, const uint64_t* lru, const uint8_t* pdev, const double* pconf,
                       cudaStream_t st);
int launch_tombstone(DeviceState& s, long long n, const int64_t* slot_dev, cudaStream_t st);
int launch_gather_rows(DeviceState& s, long long n, const int64_t* slot_dev, float* out, cudaStream_t st);
int launch_prep_queries(DeviceState& s, int m, const float* q_dev, const uint64_t* q_ctx_src, cudaStream_t st,
                        const uint32_t* wait_flags = nullptr, int n_wait = 0, uint32_t wait_epoch = 0);
int launch_scatter_inputs(int world, int q0, int q1, int dim, const float* src_q, const uint64_t* src_ctx,
                          float* const* dst_q, uint64_t* const* dst_ctx, uint32_t* const* flag, uint32_t epoch,
                          unsigned* ticket, cudaStream_t st);
int launch_prep_centroids(DeviceState& s, int k, cudaStream_t st);
int launch_gemm_topk(DeviceState& s, int m, long long n_rows, const qre_params_t& p,
                     float* dump, long long dump_ld, long long dump_row0, int dump_n, cudaStream_t st);
int launch_gemm_topk_pair(DeviceState& s, int m, long long n_rows, const qre_params_t& p,
                          float* dump, long long dump_ld, long long dump_row0, cudaStream_t st);
// Shared by both scan launchers: the argument block of one launch on the current slot.
GemmArgs make_gemm_args(const DeviceState& s, const ScanGeometry& g, int m, long long n_rows, const qre_params_t& p,
                        float* dump, long long dump_ld, long long dump_row0);
int launch_rescore(DeviceState& s, int m, long long n_rows, const qre_params_t& p, int shard,
                   qre_candidate_t* cand, cudaStream_t st);
int launch_exact_scan(DeviceState& s, int m, long long n_rows, const int32_t* qlist_dev,
                      int nq_total, bool use_ovf_list, const qre_params_t& p, int shard, qre_candidate_t* cand,
                      cudaStream_t st);
int launch_overflow_settle(DeviceState& s, long long n_rows, const qre_params_t& p, int shard,
                           qre_candidate_t* cand, cudaStream_t st);
int launch_merge_route(DeviceState& s, int m, int n_shards, const qre_candidate_t* cand_all, const int32_t* tok,
                       const uint8_t* heur_dev, const double* heur_conf, const qre_params_t& p,
                       const qre_result_t& out, cudaStream_t st, long long shard_stride = 0,
                       const uint32_t* wait_flags = nullptr, uint32_t wait_epoch = 0, int q0 = -1, int q1 = -1);
// The routing half of merge_route alone (needs only the prepared queries): the pipeline runs it
// beside the scan instead of after it.
int launch_route_only(DeviceState& s, int m, const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                      const qre_params_t& p, const qre_result_t& out, cudaStream_t st, int q0 = -1, int q1 = -1);
int launch_publish(int m, int world, const qre_candidate_t* cand, qre_candidate_t* const* dst, uint32_t* const* flag,
                   uint32_t epoch, unsigned* ticket, cudaStream_t st, int per = 0);

}  // namespace qre
