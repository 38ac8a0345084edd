// qre_api.cu — the C ABI of libqre_b200.so (declared in include/qre_b200.h): handle
// lifetime, host<->device staging, TMA descriptor set-up and the launch sequences.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "qre_common.cuh"
#include "qre_encoder.cuh"

using namespace qre;

namespace {

thread_local std::string g_err;

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define QRE_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e__ = (expr);                                                                   \
    if (e__ != cudaSuccess)                                                                     \
      return fail(QRE_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// The driver API is reached through the runtime so the library has no link-time dependency
// on libcuda.so (it must load, and export its symbols, on a machine without a driver).
int get_encode_fn(EncodeTiledFn* fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  QRE_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
  if (p == nullptr || qres != cudaDriverEntryPointSuccess)
    return fail(QRE_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
  *fn = reinterpret_cast<EncodeTiledFn>(p);
  return QRE_OK;
}

int make_tmap_bf16(EncodeTiledFn enc, CUtensorMap* map, void* base, uint64_t inner, uint64_t rows,
                   uint32_t box_inner, uint32_t box_rows) {
  cuuint64_t dims[2] = {inner, rows};
  cuuint64_t strides[1] = {inner * 2};
  cuuint32_t box[2] = {box_inner, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(QRE_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return QRE_OK;
}

struct Section {
  size_t off = 0, bytes = 0;
};

// Device-side result block (one allocation, one D2H copy).
struct ResultLayout {
  Section hit_slot, hit_sim, hit_pred_dev, hit_pred_conf, hit_shard, cand, centroid_sim, tok_dev, tok_conf,
      sem_dev, sem_conf, sem_method, dev, conf, score, flags, ovf;
  size_t total = 0;
  void plan(int m, int K, int D) {
    size_t o = 0;
    auto add = [&](Section& s, size_t bytes) {
      s.off = o;
      s.bytes = bytes;
      o = align_up(o + bytes, 256);
    };
    add(hit_slot, sizeof(int64_t) * m);
    add(hit_sim, sizeof(float) * m);
    add(hit_pred_dev, m);
    add(hit_pred_conf, sizeof(double) * m);
    add(hit_shard, sizeof(uint32_t) * m);
    add(cand, sizeof(qre_candidate_t) * m * QRE_TOPK);
    add(centroid_sim, sizeof(float) * m * (K > 0 ? K : 1));
    add(tok_dev, m);
    add(tok_conf, sizeof(double) * m);
    add(sem_dev, m);
    add(sem_conf, sizeof(double) * m);
    add(sem_method, m);
    add(dev, m);
    add(conf, sizeof(double) * m);
    add(score, sizeof(double) * m * D);
    add(flags, m);
    add(ovf, sizeof(int32_t));
    total = o;
  }
  qre_result_t view(uint8_t* base) const {
    qre_result_t r;
    r.hit_slot = reinterpret_cast<int64_t*>(base + hit_slot.off);
    r.hit_sim = reinterpret_cast<float*>(base + hit_sim.off);
    r.hit_pred_dev = base + hit_pred_dev.off;
    r.hit_pred_conf = reinterpret_cast<double*>(base + hit_pred_conf.off);
    r.hit_shard = reinterpret_cast<uint32_t*>(base + hit_shard.off);
    r.cand = reinterpret_cast<qre_candidate_t*>(base + cand.off);
    r.centroid_sim = reinterpret_cast<float*>(base + centroid_sim.off);
    r.tok_dev = base + tok_dev.off;
    r.tok_conf = reinterpret_cast<double*>(base + tok_conf.off);
    r.sem_dev = base + sem_dev.off;
    r.sem_conf = reinterpret_cast<double*>(base + sem_conf.off);
    r.sem_method = base + sem_method.off;
    r.dev = base + dev.off;
    r.conf = reinterpret_cast<double*>(base + conf.off);
    r.score = reinterpret_cast<double*>(base + score.off);
    r.flags = base + flags.off;
    return r;
  }
};

constexpr size_t kStageBytes = 64u << 20;   // device staging for host-side append/update

}  // namespace

// Helpers shared with the other translation units of the library (qre_encoder.cuh).
namespace qre {
int set_error(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
int make_tensor_map_bf16(CUtensorMap* map, void* base, uint64_t inner, uint64_t rows, uint32_t box_inner,
                         uint32_t box_rows) {
  EncodeTiledFn enc = nullptr;
  int rc = get_encode_fn(&enc);
  if (rc != QRE_OK) return rc;
  return make_tmap_bf16(enc, map, base, inner, rows, box_inner, box_rows);
}
}  // namespace qre

struct qre_cache {
  DeviceState s;
  long long hwm = 0, live = 0;
  std::vector<int64_t> free_slots;
  std::vector<uint8_t> slot_live;
  uint8_t* d_stage = nullptr;          // kStageBytes
  // Two host-API slots: batch i+1's inputs can be copied in while batch i is being scanned.
  struct HostSlot {
    uint8_t* d_in = nullptr;           // lookup inputs (H2D target)
    uint8_t* h_in = nullptr;           // pinned staging
    uint8_t* d_res = nullptr;          // result block
    uint8_t* h_res = nullptr;          // pinned
    cudaEvent_t e_in = nullptr, e_done = nullptr;
    bool pending = false;
    int m = 0, q0 = 0, q1 = 0;           // [q0, q1): queries whose results this call returns
    qre_params_t p;
    ResultLayout L;
    const int32_t* d_tok = nullptr;
    const uint8_t* d_hd = nullptr;
    const double* d_hc = nullptr;
  } hs[kApiSlots];
  cudaStream_t h2d = nullptr, d2h = nullptr;
  // The pipeline (qre_submit_dev / qre_lookup_route_begin): query preparation on `aux`, the scan
  // alone on the handle's main stream, re-score / settle / exchange / merge + route on `post`.
  cudaStream_t aux = nullptr, post = nullptr;
  // Consecutive large-batch scans alternate between the main stream and this one (same priority):
  // as two independent kernels the next scan's CTAs move onto SMs as the previous scan's CTAs retire,
  // so its set-up and the previous scan's tail overlap instead of adding up (QRE_ONE_SCAN_STREAM=1: off).
  cudaStream_t scan2 = nullptr;
  bool alt_scan = true;
  cudaEvent_t ev_in[kApiSlots] = {};                                 // per API slot: inputs valid
  cudaEvent_t ev_prep[kPipeSlots] = {}, ev_route[kPipeSlots] = {}, ev_scan[kPipeSlots] = {}, ev_post[kPipeSlots] = {};
  bool post_used[kPipeSlots] = {};
  uint64_t pipe_seq = 0;               // batches enqueued so far; batch i uses pipeline slot i % kPipeSlots
  int api_ps[kApiSlots] = {-1, -1, -1};  // pipeline slot of the batch last submitted on each API slot
  size_t in_bytes = 0;
  size_t res_bytes = 0;
  int last_slot = 0;
  qre_candidate_t* d_gather = nullptr;  // qre_group_lookup_route: [16][max_batch * QRE_TOPK] lists of the group
  cudaEvent_t ev_group = nullptr;
  float* d_dump = nullptr;
  size_t dump_bytes = 0;
  int64_t launches = 0;
  int last_m = 0;
  // peer exchange (qre_xchg_*): [2 slots][world][max_batch * QRE_TOPK] records, then
  // [2][world][32] uint32 flags; mapped into every peer process with CUDA IPC
  uint8_t* xchg = nullptr;
  size_t xchg_flags_off = 0;
  // input exchange of the host-buffer path: [kPipeSlots]{queries fp32 [max_batch, dim], ctx [max_batch]},
  // per-rank arrival flags [kPipeSlots][world][32]; same allocation, same offsets on every rank
  size_t xin_flags_off = 0, xin_off = 0, xin_slot_bytes = 0, xin_ctx_off = 0;
  uint32_t xin_epoch = 0;
  unsigned* xin_ticket = nullptr;
  int xchg_world = 0, xchg_rank = 0;
  uint8_t* xchg_peer[16] = {};
  bool xchg_open[16] = {};
  uint32_t xchg_pub_epoch = 0, xchg_merge_epoch = 0;
  unsigned* xchg_ticket = nullptr;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
  }
  ~DeviceGuard() {
    int cur = -1;
    cudaGetDevice(&cur);
    if (prev >= 0 && cur != prev) cudaSetDevice(prev);
  }
};

size_t input_bytes(int m, int dim) {
  return align_up(sizeof(float) * (size_t)m * dim, 256) + align_up(sizeof(uint64_t) * (size_t)m, 256) +
         align_up(sizeof(int32_t) * (size_t)m, 256) + align_up((size_t)m, 256) +
         align_up(sizeof(double) * (size_t)m, 256);
}

int validate_params(const qre_cache_t* h, const qre_params_t* p) {
  if (p == nullptr) return fail(QRE_ERR_INVALID, "params is NULL");
  if (p->scan_mode != QRE_SCAN_TENSOR && p->scan_mode != QRE_SCAN_EXACT)
    return fail(QRE_ERR_INVALID, "unknown scan_mode %d", p->scan_mode);
  if (p->do_route) {
    if (p->n_devices < 1 || p->n_devices > QRE_MAX_DEVICES)
      return fail(QRE_ERR_INVALID, "n_devices %d out of range [1, %d]", p->n_devices, QRE_MAX_DEVICES);
    if (h->s.n_centroids < 1) return fail(QRE_ERR_INVALID, "do_route set but no centroids uploaded");
  }
  return QRE_OK;
}

// After the scan: exact re-score and, when the spill pool ran out, the device-driven brute-force pass.
int rescore_enqueue(qre_cache_t* h, int m, const qre_params_t& p, int shard, qre_candidate_t* cand,
                    cudaStream_t st) {
  DeviceState& s = h->s;
  if (p.scan_mode == QRE_SCAN_TENSOR && h->hwm > 0) {
    h->launches += launch_rescore(s, m, h->hwm, p, shard, cand, st);
    h->launches += launch_overflow_settle(s, h->hwm, p, shard, cand, st);
  } else {
    h->launches += launch_exact_scan(s, m, h->hwm, nullptr, m, false, p, shard, cand, st);
  }
  QRE_CUDA(cudaGetLastError());
  return QRE_OK;
}

int tensor_scan_enqueue(qre_cache_t* h, int m, const qre_params_t& p, cudaStream_t st) {
  if (p.scan_mode != QRE_SCAN_TENSOR || h->hwm <= 0) return QRE_OK;
  const int n = launch_gemm_topk(h->s, m, h->hwm, p, nullptr, 0, 0, 0, st);
  if (n < 0) return fail(QRE_ERR_CAPACITY, "scan geometry for m=%d, %lld rows does not fit the top-k scratch", m,
                         (long long)h->hwm);
  h->launches += n;
  QRE_CUDA(cudaGetLastError());
  return QRE_OK;
}

// queries -> this shard's candidates (device pointers, enqueue only), everything on one stream
int scan_enqueue(qre_cache_t* h, int m, const float* q_dev, const uint64_t* q_ctx_dev, const qre_params_t& p,
                 int shard, qre_candidate_t* cand, cudaStream_t st) {
  DeviceState& s = h->s;
  s.slot = p.pipeline_slot & 1;
  h->last_slot = s.slot;
  h->launches += launch_prep_queries(s, m, q_dev, q_ctx_dev, st);
  h->last_m = m;
  int rc = tensor_scan_enqueue(h, m, p, st);
  if (rc != QRE_OK) return rc;
  return rescore_enqueue(h, m, p, shard, cand, st);
}

}  // namespace

extern "C" {

int qre_abi_version(void) { return QRE_ABI_VERSION; }

const char* qre_last_error(void) { return g_err.c_str(); }

int qre_device_count(int* count) {
  if (count == nullptr) return fail(QRE_ERR_INVALID, "count is NULL");
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess) {
    *count = 0;
    return fail(QRE_ERR_NO_DEVICE, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  }
  *count = n;
  return QRE_OK;
}

int qre_cache_create(int device, int dim, int64_t capacity, int max_batch, qre_cache_t** out) {
  if (out == nullptr) return fail(QRE_ERR_INVALID, "out is NULL");
  *out = nullptr;
  if (dim < 8 || dim % 8 != 0) return fail(QRE_ERR_INVALID, "dim must be a positive multiple of 8 (got %d)", dim);
  // row numbers travel as int32 TMA coordinates and u32 candidate slots
  if (capacity < 1 || capacity > 2147483000ll) return fail(QRE_ERR_INVALID, "capacity out of range [1, 2^31)");
  if (max_batch < 1 || max_batch > (1 << 20)) return fail(QRE_ERR_INVALID, "max_batch out of range");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1)
    return fail(QRE_ERR_NO_DEVICE, "no CUDA device: this library has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(QRE_ERR_INVALID, "device %d out of range (have %d)", device, ndev);
  cudaDeviceProp prop;
  QRE_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(QRE_ERR_NO_DEVICE, "device %d is sm_%d%d; libqre_b200 is built for sm_100a only", device,
                prop.major, prop.minor);
  const int dim_pad = (int)align_up(dim, kBK);
  if (dim_pad / kBK > kMaxKBStream)
    return fail(QRE_ERR_INVALID, "dim %d not supported (max %d)", dim, kMaxKBStream * kBK);
  DeviceGuard guard(device);
  qre_cache_t* h = new (std::nothrow) qre_cache_t();
  if (h == nullptr) return fail(QRE_ERR_INVALID, "out of host memory");
  DeviceState& s = h->s;
  s.device = device;
  s.num_sms = prop.multiProcessorCount;
  s.dim = dim;
  s.dim_pad = dim_pad;
  s.max_batch = max_batch;
  s.mb_pad = (int)align_up(max_batch, 2 * kBM);     // CTA pairs work on two 128-query blocks
  s.capacity = capacity;
  s.cap_pad = (long long)align_up((size_t)capacity, kBN);
#define QRE_TRY(expr)                 \
  do {                                \
    int rc__ = (expr);                \
    if (rc__ != QRE_OK) {             \
      std::string keep = g_err;       \
      qre_cache_destroy(h);           \
      g_err = keep;                   \
      return rc__;                    \
    }                                 \
  } while (0)
#define QRE_ALLOC(ptr, bytes) QRE_TRY([&]() -> int { QRE_CUDA(cudaMalloc((void**)&(ptr), (bytes))); return QRE_OK; }())
  // The scan's stream gets the highest priority, the streams of the small kernels that run beside it
  // the lowest: when an SM frees up between two scans the block scheduler places the next scan's
  // CTA (which needs nearly all of the SM's shared memory) before any queued small block.
  int prio_least = 0, prio_greatest = 0;
  QRE_TRY([&]() -> int {
    QRE_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    QRE_CUDA(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, prio_greatest));
    return QRE_OK;
  }());
  QRE_ALLOC(s.rows_f32, sizeof(float) * (size_t)capacity * dim_pad);
  QRE_ALLOC(s.rows_bf16, sizeof(__nv_bfloat16) * (size_t)s.cap_pad * dim_pad);
  QRE_ALLOC(s.row_norm, sizeof(float) * (size_t)capacity);
  QRE_ALLOC(s.meta, sizeof(RowMeta) * (size_t)capacity);
  QRE_ALLOC(s.lru, sizeof(uint64_t) * (size_t)capacity);
  QRE_ALLOC(s.pred_dev, (size_t)capacity);
  QRE_ALLOC(s.pred_conf, sizeof(double) * (size_t)capacity);
  // top-k scratch from the hard bound on work units: at most 8 rounds of the SM count (or one unit
  // per query block when the batch alone exceeds that), 128 queries x 8 entries per unit
  s.partial_entries = 1024LL * (8LL * s.num_sms + s.mb_pad / kBM + 2);
  {
    const long long want = std::max<long long>(1024, 128LL * max_batch);     // ~1,900 chained candidates / query
    s.spill_cap = (uint32_t)std::min<long long>(want, 1LL << 22);            // <= 512 MB
  }
  for (int b = 0; b < kPipeSlots; ++b) {
    DeviceState::Batch& B = s.batch[b];
    QRE_ALLOC(B.q_f32, sizeof(float) * (size_t)s.mb_pad * dim_pad);
    QRE_ALLOC(B.q_bf16, sizeof(__nv_bfloat16) * (size_t)s.mb_pad * dim_pad);
    QRE_ALLOC(B.q_norm, sizeof(float) * (size_t)s.mb_pad);
    QRE_ALLOC(B.q_ctx, sizeof(uint64_t) * (size_t)max_batch);
    QRE_ALLOC(B.q_flags, (size_t)max_batch);
    QRE_ALLOC(B.partial, sizeof(PartialEntry) * (size_t)s.partial_entries);
    QRE_ALLOC(B.spill_head, sizeof(uint32_t) * (size_t)(s.partial_entries / kTopK));
    QRE_ALLOC(B.spill_pool, sizeof(SpillChunk) * (size_t)s.spill_cap);
    QRE_ALLOC(B.counters, sizeof(uint32_t) * 8);
    QRE_TRY([&]() -> int { QRE_CUDA(cudaMemset(B.counters, 0, sizeof(uint32_t) * 8)); return QRE_OK; }());
    QRE_ALLOC(B.ovf_list, sizeof(int32_t) * (size_t)max_batch);
    QRE_ALLOC(B.cand, sizeof(qre_candidate_t) * (size_t)max_batch * QRE_TOPK);
  }
  s.exact_grid = s.num_sms * 2;
  QRE_ALLOC(s.exact_partial, sizeof(ExactBest) * (size_t)kExactGroup * s.exact_grid);
  s.settle_groups = (max_batch + kExactGroup - 1) / kExactGroup;
  QRE_ALLOC(s.settle_tickets, sizeof(unsigned) * (size_t)s.settle_groups);
  QRE_ALLOC(s.settle_partial, sizeof(ExactBest) * (size_t)s.settle_groups * kExactGroup * s.exact_grid);
  QRE_ALLOC(s.centroids, sizeof(float) * (size_t)QRE_MAX_CENTROIDS * dim_pad);
  QRE_ALLOC(s.cent_norm, sizeof(float) * QRE_MAX_CENTROIDS);
  QRE_ALLOC(s.cent_owner, QRE_MAX_CENTROIDS);
  QRE_ALLOC(s.perf_health, sizeof(double) * QRE_MAX_DEVICES);
  QRE_ALLOC(h->d_stage, kStageBytes);
  h->in_bytes = input_bytes(max_batch, dim);
  {
    ResultLayout L;
    L.plan(max_batch, QRE_MAX_CENTROIDS, QRE_MAX_DEVICES);
    h->res_bytes = L.total;
  }
  for (int b = 0; b < kApiSlots; ++b) {
    QRE_ALLOC(h->hs[b].d_in, h->in_bytes);
    QRE_ALLOC(h->hs[b].d_res, h->res_bytes);
  }
  QRE_TRY([&]() -> int {
    for (int b = 0; b < kApiSlots; ++b) {
      QRE_CUDA(cudaMallocHost((void**)&h->hs[b].h_in, h->in_bytes));
      QRE_CUDA(cudaMallocHost((void**)&h->hs[b].h_res, h->res_bytes));
      QRE_CUDA(cudaEventCreateWithFlags(&h->hs[b].e_in, cudaEventDisableTiming));
      QRE_CUDA(cudaEventCreateWithFlags(&h->hs[b].e_done, cudaEventDisableTiming));
      QRE_CUDA(cudaEventCreateWithFlags(&h->ev_in[b], cudaEventDisableTiming));
    }
    for (int b = 0; b < kPipeSlots; ++b) {
      QRE_CUDA(cudaEventCreateWithFlags(&h->ev_prep[b], cudaEventDisableTiming));
      QRE_CUDA(cudaEventCreateWithFlags(&h->ev_route[b], cudaEventDisableTiming));
      QRE_CUDA(cudaEventCreateWithFlags(&h->ev_scan[b], cudaEventDisableTiming));
      QRE_CUDA(cudaEventCreateWithFlags(&h->ev_post[b], cudaEventDisableTiming));
    }
    QRE_CUDA(cudaStreamCreateWithFlags(&h->h2d, cudaStreamNonBlocking));
    QRE_CUDA(cudaStreamCreateWithFlags(&h->d2h, cudaStreamNonBlocking));
    QRE_CUDA(cudaStreamCreateWithPriority(&h->aux, cudaStreamNonBlocking, prio_least));
    QRE_CUDA(cudaStreamCreateWithPriority(&h->post, cudaStreamNonBlocking, prio_least));
    QRE_CUDA(cudaStreamCreateWithPriority(&h->scan2, cudaStreamNonBlocking, prio_greatest));
    QRE_CUDA(cudaMemsetAsync(s.rows_bf16, 0, sizeof(__nv_bfloat16) * (size_t)s.cap_pad * dim_pad, s.stream));
    QRE_CUDA(cudaMemsetAsync(s.meta, 0xFF, sizeof(RowMeta) * (size_t)capacity, s.stream));
    QRE_CUDA(cudaMemsetAsync(s.row_norm, 0, sizeof(float) * (size_t)capacity, s.stream));
    QRE_CUDA(cudaMemsetAsync(s.perf_health, 0, sizeof(double) * QRE_MAX_DEVICES, s.stream));
    QRE_CUDA(cudaMemsetAsync(s.settle_tickets, 0, sizeof(unsigned) * (size_t)s.settle_groups, s.stream));
    for (int b = 0; b < kPipeSlots; ++b) {
      QRE_CUDA(cudaMemsetAsync(s.batch[b].q_bf16, 0, sizeof(__nv_bfloat16) * (size_t)s.mb_pad * dim_pad, s.stream));
      QRE_CUDA(cudaMemsetAsync(s.batch[b].counters, 0, sizeof(uint32_t) * 4, s.stream));
      QRE_CUDA(cudaMemsetAsync(s.batch[b].q_flags, 0, (size_t)max_batch, s.stream));
    }
    QRE_CUDA(gemm_topk_init());
    QRE_CUDA(gemm_topk_pair_init());
    {
      const char* e = getenv("QRE_SCAN_SINGLE_CTA");
      s.allow_pair = !(e != nullptr && e[0] == '1');
      const char* o = getenv("QRE_ONE_SCAN_STREAM");
      h->alt_scan = !(o != nullptr && o[0] == '1');
    }
    EncodeTiledFn enc = nullptr;
    int rc = get_encode_fn(&enc);
    if (rc != QRE_OK) return rc;
    rc = make_tmap_bf16(enc, &s.tmap_e, s.rows_bf16, (uint64_t)dim_pad, (uint64_t)s.cap_pad, kBK, kBN);
    if (rc != QRE_OK) return rc;
    rc = make_tmap_bf16(enc, &s.tmap_eh, s.rows_bf16, (uint64_t)dim_pad, (uint64_t)s.cap_pad, kBK, kBN / 2);
    if (rc != QRE_OK) return rc;
    for (int b = 0; b < kPipeSlots; ++b) {
      rc = make_tmap_bf16(enc, &s.batch[b].tmap_q, s.batch[b].q_bf16, (uint64_t)dim_pad, (uint64_t)s.mb_pad, kBK,
                          kBM);
      if (rc != QRE_OK) return rc;
    }
    QRE_CUDA(cudaStreamSynchronize(s.stream));
    return QRE_OK;
  }());
#undef QRE_ALLOC
#undef QRE_TRY
  h->slot_live.assign((size_t)capacity, 0);
  *out = h;
  return QRE_OK;
}

int qre_cache_destroy(qre_cache_t* h) {
  if (h == nullptr) return QRE_OK;
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  if (s.stream) cudaStreamSynchronize(s.stream);
  if (h->aux) cudaStreamSynchronize(h->aux);
  if (h->post) cudaStreamSynchronize(h->post);
  if (h->scan2) cudaStreamSynchronize(h->scan2);
  std::vector<void*> dev_ptrs = {s.rows_f32, s.rows_bf16, s.row_norm, s.meta, s.lru, s.pred_dev, s.pred_conf,
                                 s.exact_partial, s.settle_tickets, s.settle_partial, s.centroids, s.cent_norm,
                                 s.cent_owner, s.perf_health, h->d_stage, h->hs[0].d_in, h->hs[1].d_in,
                                 h->hs[2].d_in, h->hs[0].d_res, h->hs[1].d_res, h->hs[2].d_res, h->d_dump,
                                 h->d_gather};
  for (int b = 0; b < kPipeSlots; ++b) {
    DeviceState::Batch& B = s.batch[b];
    for (void* p : {(void*)B.q_f32, (void*)B.q_bf16, (void*)B.q_norm, (void*)B.q_ctx, (void*)B.q_flags,
                    (void*)B.partial, (void*)B.spill_head, (void*)B.spill_pool, (void*)B.counters,
                    (void*)B.ovf_list, (void*)B.cand})
      dev_ptrs.push_back(p);
  }
  for (void* p : dev_ptrs)
    if (p) cudaFree(p);
  for (int b = 0; b < kApiSlots; ++b) {
    if (h->hs[b].h_in) cudaFreeHost(h->hs[b].h_in);
    if (h->hs[b].h_res) cudaFreeHost(h->hs[b].h_res);
    if (h->hs[b].e_in) cudaEventDestroy(h->hs[b].e_in);
    if (h->hs[b].e_done) cudaEventDestroy(h->hs[b].e_done);
    if (h->ev_in[b]) cudaEventDestroy(h->ev_in[b]);
  }
  for (int b = 0; b < kPipeSlots; ++b)
    for (cudaEvent_t e : {h->ev_prep[b], h->ev_route[b], h->ev_scan[b], h->ev_post[b]})
      if (e) cudaEventDestroy(e);
  for (int pr = 0; pr < 16; ++pr)
    if (h->xchg_open[pr] && h->xchg_peer[pr]) cudaIpcCloseMemHandle(h->xchg_peer[pr]);
  if (h->ev_group) cudaEventDestroy(h->ev_group);
  if (h->xchg) cudaFree(h->xchg);
  if (h->xchg_ticket) cudaFree(h->xchg_ticket);
  if (h->h2d) cudaStreamDestroy(h->h2d);
  if (h->d2h) cudaStreamDestroy(h->d2h);
  if (h->aux) cudaStreamDestroy(h->aux);
  if (h->post) cudaStreamDestroy(h->post);
  if (h->scan2) cudaStreamDestroy(h->scan2);
  if (s.prof_ev) {
    for (int i = 0; i < 2 * kProfMax; ++i) cudaEventDestroy(s.prof_ev[i]);
    delete[] s.prof_ev;
  }
  if (s.stream) cudaStreamDestroy(s.stream);
  delete h;
  return QRE_OK;
}

int qre_cache_info(const qre_cache_t* h, int64_t* capacity, int64_t* high_water, int64_t* live, int* dim,
                   int* max_batch, int* device) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (capacity) *capacity = h->s.capacity;
  if (high_water) *high_water = h->hwm;
  if (live) *live = h->live;
  if (dim) *dim = h->s.dim;
  if (max_batch) *max_batch = h->s.max_batch;
  if (device) *device = h->s.device;
  return QRE_OK;
}

int qre_cache_stream(const qre_cache_t* h, void** stream) {
  if (h == nullptr || stream == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  *stream = (void*)h->s.stream;
  return QRE_OK;
}

int qre_launch_count(const qre_cache_t* h, int64_t* count) {
  if (h == nullptr || count == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  *count = h->launches;
  return QRE_OK;
}

// Choose n slots: tombstoned ones first (LIFO), then fresh ones past the high-water mark.
static int take_slots(qre_cache_t* h, int64_t n, int64_t* slot_out) {
  const long long avail = (long long)h->free_slots.size() + (h->s.capacity - h->hwm);
  if (n > avail) return fail(QRE_ERR_CAPACITY, "cache shard full: need %lld slots, %lld free", (long long)n, avail);
  for (int64_t i = 0; i < n; ++i) {
    int64_t sl;
    if (!h->free_slots.empty()) {
      sl = h->free_slots.back();
      h->free_slots.pop_back();
    } else {
      sl = h->hwm++;
    }
    h->slot_live[(size_t)sl] = 1;
    slot_out[i] = sl;
  }
  h->live += n;
  return QRE_OK;
}

// One staged chunk: copy host arrays (any may be NULL) into d_stage and launch write_rows.
static int staged_write(qre_cache_t* h, int64_t n, const int64_t* slot, const float* rows, const uint64_t* ctx,
                        const int64_t* ts, const uint64_t* lru, const uint8_t* pdev, const double* pconf) {
  DeviceState& s = h->s;
  const size_t per_row = sizeof(float) * s.dim + 8 + 8 + 8 + 8 + 1 + 8;
  const int64_t chunk = std::max<int64_t>(1, (int64_t)((kStageBytes - 8 * 256) / per_row));
  for (int64_t i0 = 0; i0 < n; i0 += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - i0);
    size_t o = 0;
    auto put = [&](const void* src, size_t bytes, void** dst) -> cudaError_t {
      *dst = nullptr;
      if (src == nullptr) return cudaSuccess;
      *dst = h->d_stage + o;
      o = align_up(o + bytes, 256);
      return cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, s.stream);
    };
    void *d_slot, *d_rows, *d_ctx, *d_ts, *d_lru, *d_pdev, *d_pconf;
    QRE_CUDA(put(slot + i0, sizeof(int64_t) * c, &d_slot));
    QRE_CUDA(put(rows ? rows + i0 * s.dim : nullptr, sizeof(float) * c * s.dim, &d_rows));
    QRE_CUDA(put(ctx ? ctx + i0 : nullptr, sizeof(uint64_t) * c, &d_ctx));
    QRE_CUDA(put(ts ? ts + i0 : nullptr, sizeof(int64_t) * c, &d_ts));
    QRE_CUDA(put(lru ? lru + i0 : nullptr, sizeof(uint64_t) * c, &d_lru));
    QRE_CUDA(put(pdev ? pdev + i0 : nullptr, (size_t)c, &d_pdev));
    QRE_CUDA(put(pconf ? pconf + i0 : nullptr, sizeof(double) * c, &d_pconf));
    h->launches += launch_append(s, c, (const float*)d_rows, s.dim, (const int64_t*)d_slot, (const uint64_t*)d_ctx,
                                 (const int64_t*)d_ts, (const uint64_t*)d_lru, (const uint8_t*)d_pdev,
                                 (const double*)d_pconf, s.stream);
    QRE_CUDA(cudaGetLastError());
    // pageable sources: the copies above have consumed the host data before returning, and the
    // next chunk's copies are stream-ordered after this kernel, so d_stage can be reused.
  }
  QRE_CUDA(cudaStreamSynchronize(s.stream));
  return QRE_OK;
}

int qre_cache_append(qre_cache_t* h, int64_t n, const float* rows, const uint64_t* ctx, const int64_t* ts_us,
                     const uint64_t* lru_seq, const uint8_t* pred_dev, const double* pred_conf,
                     int64_t* slot_out) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n < 0 || (n > 0 && (!rows || !ctx || !ts_us || !lru_seq || !pred_dev || !pred_conf || !slot_out)))
    return fail(QRE_ERR_INVALID, "append needs every array");
  if (n == 0) return QRE_OK;
  DeviceGuard guard(h->s.device);
  int rc = take_slots(h, n, slot_out);
  if (rc != QRE_OK) return rc;
  return staged_write(h, n, slot_out, rows, ctx, ts_us, lru_seq, pred_dev, pred_conf);
}

int qre_cache_append_dev(qre_cache_t* h, int64_t n, const float* rows, const uint64_t* ctx, const int64_t* ts_us,
                         const uint64_t* lru_seq, const uint8_t* pred_dev, const double* pred_conf,
                         int64_t* slot_out_host, void* stream) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n < 0 || (n > 0 && (!rows || !ctx || !ts_us || !lru_seq || !pred_dev || !pred_conf)))
    return fail(QRE_ERR_INVALID, "append_dev needs every array");
  if (n == 0) return QRE_OK;
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  cudaStream_t st = stream ? (cudaStream_t)stream : s.stream;
  std::vector<int64_t> slots((size_t)n);
  int rc = take_slots(h, n, slots.data());
  if (rc != QRE_OK) return rc;
  if (slot_out_host) memcpy(slot_out_host, slots.data(), sizeof(int64_t) * (size_t)n);
  const int64_t chunk = (int64_t)(kStageBytes / sizeof(int64_t));
  for (int64_t i0 = 0; i0 < n; i0 += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - i0);
    QRE_CUDA(cudaMemcpyAsync(h->d_stage, slots.data() + i0, sizeof(int64_t) * c, cudaMemcpyHostToDevice, st));
    h->launches += launch_append(s, c, rows + i0 * s.dim, s.dim, (const int64_t*)h->d_stage, ctx + i0, ts_us + i0,
                                 lru_seq + i0, pred_dev + i0, pred_conf + i0, st);
    QRE_CUDA(cudaGetLastError());
    QRE_CUDA(cudaStreamSynchronize(st));   // `slots` is pageable and d_stage is reused
  }
  return QRE_OK;
}

static int check_slots(const qre_cache_t* h, int64_t n, const int64_t* slot) {
  for (int64_t i = 0; i < n; ++i)
    if (slot[i] < 0 || slot[i] >= h->hwm || !h->slot_live[(size_t)slot[i]])
      return fail(QRE_ERR_INVALID, "slot %lld is not a live row", (long long)slot[i]);
  return QRE_OK;
}

int qre_cache_update(qre_cache_t* h, int64_t n, const int64_t* slot, const float* rows, const uint64_t* ctx,
                     const int64_t* ts_us, const uint64_t* lru_seq, const uint8_t* pred_dev,
                     const double* pred_conf) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n < 0 || (n > 0 && !slot)) return fail(QRE_ERR_INVALID, "slot is NULL");
  if (n == 0) return QRE_OK;
  int rc = check_slots(h, n, slot);
  if (rc != QRE_OK) return rc;
  DeviceGuard guard(h->s.device);
  return staged_write(h, n, slot, rows, ctx, ts_us, lru_seq, pred_dev, pred_conf);
}

int qre_cache_tombstone(qre_cache_t* h, int64_t n, const int64_t* slot) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n < 0 || (n > 0 && !slot)) return fail(QRE_ERR_INVALID, "slot is NULL");
  if (n == 0) return QRE_OK;
  int rc = check_slots(h, n, slot);
  if (rc != QRE_OK) return rc;
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  const int64_t chunk = (int64_t)(kStageBytes / sizeof(int64_t));
  for (int64_t i0 = 0; i0 < n; i0 += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - i0);
    QRE_CUDA(cudaMemcpyAsync(h->d_stage, slot + i0, sizeof(int64_t) * c, cudaMemcpyHostToDevice, s.stream));
    h->launches += launch_tombstone(s, c, (const int64_t*)h->d_stage, s.stream);
    QRE_CUDA(cudaGetLastError());
  }
  QRE_CUDA(cudaStreamSynchronize(s.stream));
  for (int64_t i = 0; i < n; ++i) {
    if (h->slot_live[(size_t)slot[i]]) {      // a slot listed twice is freed once
      h->slot_live[(size_t)slot[i]] = 0;
      h->free_slots.push_back(slot[i]);
      --h->live;
    }
  }
  return QRE_OK;
}

int qre_cache_clear(qre_cache_t* h) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  QRE_CUDA(cudaMemsetAsync(s.rows_bf16, 0, sizeof(__nv_bfloat16) * (size_t)s.cap_pad * s.dim_pad, s.stream));
  QRE_CUDA(cudaMemsetAsync(s.meta, 0xFF, sizeof(RowMeta) * (size_t)s.capacity, s.stream));
  QRE_CUDA(cudaMemsetAsync(s.row_norm, 0, sizeof(float) * (size_t)s.capacity, s.stream));
  QRE_CUDA(cudaStreamSynchronize(s.stream));
  h->hwm = 0;
  h->live = 0;
  h->free_slots.clear();
  std::fill(h->slot_live.begin(), h->slot_live.end(), 0);
  return QRE_OK;
}

int qre_cache_read_rows(qre_cache_t* h, int64_t n, const int64_t* slot, float* rows_out) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n < 0 || (n > 0 && (!slot || !rows_out))) return fail(QRE_ERR_INVALID, "NULL argument");
  if (n == 0) return QRE_OK;
  int rc = check_slots(h, n, slot);
  if (rc != QRE_OK) return rc;
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  const size_t per_row = sizeof(int64_t) + sizeof(float) * s.dim;
  const int64_t chunk = std::max<int64_t>(1, (int64_t)((kStageBytes - 512) / per_row));
  for (int64_t i0 = 0; i0 < n; i0 += chunk) {
    const int64_t c = std::min<int64_t>(chunk, n - i0);
    int64_t* d_slot = (int64_t*)h->d_stage;
    float* d_out = (float*)(h->d_stage + align_up(sizeof(int64_t) * c, 256));
    QRE_CUDA(cudaMemcpyAsync(d_slot, slot + i0, sizeof(int64_t) * c, cudaMemcpyHostToDevice, s.stream));
    h->launches += launch_gather_rows(s, c, d_slot, d_out, s.stream);
    QRE_CUDA(cudaGetLastError());
    QRE_CUDA(cudaMemcpyAsync(rows_out + i0 * s.dim, d_out, sizeof(float) * c * s.dim, cudaMemcpyDeviceToHost,
                             s.stream));
    QRE_CUDA(cudaStreamSynchronize(s.stream));
  }
  return QRE_OK;
}

int qre_set_centroids(qre_cache_t* h, int n_centroids, const float* centroids, const uint8_t* owner_device) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n_centroids < 1 || n_centroids > QRE_MAX_CENTROIDS || !centroids || !owner_device)
    return fail(QRE_ERR_INVALID, "n_centroids out of range [1, %d] or NULL array", QRE_MAX_CENTROIDS);
  for (int k = 0; k < n_centroids; ++k)
    if (owner_device[k] >= QRE_MAX_DEVICES) return fail(QRE_ERR_INVALID, "centroid %d: owner out of range", k);
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  QRE_CUDA(cudaMemsetAsync(s.centroids, 0, sizeof(float) * (size_t)n_centroids * s.dim_pad, s.stream));
  QRE_CUDA(cudaMemcpy2DAsync(s.centroids, sizeof(float) * s.dim_pad, centroids, sizeof(float) * s.dim,
                             sizeof(float) * s.dim, n_centroids, cudaMemcpyHostToDevice, s.stream));
  QRE_CUDA(cudaMemcpyAsync(s.cent_owner, owner_device, n_centroids, cudaMemcpyHostToDevice, s.stream));
  s.n_centroids = n_centroids;
  h->launches += launch_prep_centroids(s, n_centroids, s.stream);
  QRE_CUDA(cudaGetLastError());
  QRE_CUDA(cudaStreamSynchronize(s.stream));
  return QRE_OK;
}

int qre_set_perf_health(qre_cache_t* h, int n_devices, const double* health) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (n_devices < 1 || n_devices > QRE_MAX_DEVICES || !health) return fail(QRE_ERR_INVALID, "bad n_devices");
  DeviceGuard guard(h->s.device);
  QRE_CUDA(cudaMemcpyAsync(h->s.perf_health, health, sizeof(double) * n_devices, cudaMemcpyHostToDevice,
                           h->s.stream));
  QRE_CUDA(cudaStreamSynchronize(h->s.stream));
  return QRE_OK;
}

int qre_scan_dev(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx, const qre_params_t* params,
                 int shard, qre_candidate_t* cand_out, int32_t* overflow_count, void* stream) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  if (!queries || !q_ctx || !cand_out) return fail(QRE_ERR_INVALID, "NULL argument");
  if (params == nullptr) return fail(QRE_ERR_INVALID, "params is NULL");
  if (params->scan_mode != QRE_SCAN_TENSOR && params->scan_mode != QRE_SCAN_EXACT)
    return fail(QRE_ERR_INVALID, "unknown scan_mode %d", params->scan_mode);
  DeviceGuard guard(h->s.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : h->s.stream;
  int rc = scan_enqueue(h, m, queries, q_ctx, *params, shard, cand_out, st);
  if (rc != QRE_OK) return rc;
  if (overflow_count)
    QRE_CUDA(cudaMemcpyAsync(overflow_count, h->s.cur().ovf_unresolved(), sizeof(int32_t), cudaMemcpyDeviceToDevice,
                             st));
  return QRE_OK;
}

int qre_resolve_overflow_dev(qre_cache_t* h, int m, const qre_params_t* params, int shard,
                             qre_candidate_t* cand_inout, void* stream) {
  if (h == nullptr || params == nullptr || cand_inout == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (m != h->last_m) return fail(QRE_ERR_INVALID, "m=%d does not match the last scan (%d)", m, h->last_m);
  DeviceGuard guard(h->s.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : h->s.stream;
  h->s.slot = params->pipeline_slot & 1;
  const DeviceState::Batch& B = h->s.cur();
  int32_t cnt[2] = {0, 0};
  QRE_CUDA(cudaMemcpyAsync(&cnt[0], B.ovf_count(), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  QRE_CUDA(cudaMemcpyAsync(&cnt[1], B.ovf_unresolved(), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  QRE_CUDA(cudaStreamSynchronize(st));
  const int32_t n_ovf = cnt[0];
  if (cnt[1] <= 0 || n_ovf <= 0) return QRE_OK;
  h->launches += launch_exact_scan(h->s, m, h->hwm, B.ovf_list, n_ovf, true, *params, shard, cand_inout, st);
  QRE_CUDA(cudaGetLastError());
  QRE_CUDA(cudaMemsetAsync(B.ovf_unresolved(), 0, sizeof(int32_t), st));
  return QRE_OK;
}

int qre_merge_route_dev(qre_cache_t* h, int m, int n_shards, const qre_candidate_t* cand_all, const int32_t* tok,
                        const uint8_t* heur_dev, const double* heur_conf, const qre_params_t* params,
                        qre_result_t* out_dev, void* stream) {
  if (h == nullptr || cand_all == nullptr || out_dev == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  if (n_shards < 1 || n_shards * QRE_TOPK > 128) return fail(QRE_ERR_INVALID, "n_shards out of range [1, 16]");
  int rc = validate_params(h, params);
  if (rc != QRE_OK) return rc;
  if (params->do_route && (!tok || !heur_dev || !heur_conf))
    return fail(QRE_ERR_INVALID, "do_route needs tok, heur_dev and heur_conf");
  DeviceGuard guard(h->s.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : h->s.stream;
  h->s.slot = params->pipeline_slot & 1;
  h->launches += launch_merge_route(h->s, m, n_shards, cand_all, tok, heur_dev, heur_conf, *params, *out_dev, st);
  QRE_CUDA(cudaGetLastError());
  return QRE_OK;
}

static bool is_pinned_host(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

static void scatter_results(const qre_cache_t::HostSlot& S, int K, int D, qre_result_t* out) {
  const size_t q0 = (size_t)S.q0, n = (size_t)(S.q1 - S.q0);      // sliced results: only this rank's share
  const qre_result_t hres = S.L.view(S.h_res);
#define QRE_OUT(field, elem_bytes)                                                                 \
  if (out->field && n)                                                                             \
  memcpy(reinterpret_cast<uint8_t*>(out->field) + q0 * (elem_bytes),                               \
         reinterpret_cast<const uint8_t*>(hres.field) + q0 * (elem_bytes), n * (elem_bytes))
  QRE_OUT(hit_slot, sizeof(int64_t));
  QRE_OUT(hit_sim, sizeof(float));
  QRE_OUT(hit_pred_dev, 1);
  QRE_OUT(hit_pred_conf, sizeof(double));
  QRE_OUT(hit_shard, sizeof(uint32_t));
  QRE_OUT(cand, sizeof(qre_candidate_t) * QRE_TOPK);
  QRE_OUT(flags, 1);
  if (S.p.do_route) {
    QRE_OUT(centroid_sim, sizeof(float) * (size_t)K);
    QRE_OUT(tok_dev, 1);
    QRE_OUT(tok_conf, sizeof(double));
    QRE_OUT(sem_dev, 1);
    QRE_OUT(sem_conf, sizeof(double));
    QRE_OUT(sem_method, 1);
    QRE_OUT(dev, 1);
    QRE_OUT(conf, sizeof(double));
    QRE_OUT(score, sizeof(double) * (size_t)D);
  }
#undef QRE_OUT
}

// ---- peer exchange of candidate lists (replaces the NCCL all-gather) ------------------------
static size_t xchg_list_records(const qre_cache_t* h) { return (size_t)h->s.max_batch * QRE_TOPK; }

int qre_xchg_init(qre_cache_t* h, int world, int rank, uint8_t* handle_out) {
  if (h == nullptr || handle_out == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (world < 1 || world > 16 || rank < 0 || rank >= world) return fail(QRE_ERR_INVALID, "bad world/rank");
  if (h->xchg != nullptr) return fail(QRE_ERR_INVALID, "exchange already initialised");
  static_assert(sizeof(cudaIpcMemHandle_t) == QRE_IPC_HANDLE_BYTES, "handle size");
  DeviceGuard guard(h->s.device);
  const size_t lists = sizeof(qre_candidate_t) * 2 * (size_t)world * xchg_list_records(h);
  h->xchg_flags_off = align_up(lists, 256);
  const size_t flag_bytes = sizeof(uint32_t) * 2 * (size_t)world * 32;
  const size_t xin_flag_bytes = sizeof(uint32_t) * kPipeSlots * (size_t)world * 32;
  h->xin_flags_off = h->xchg_flags_off + flag_bytes;
  h->xin_off = align_up(h->xin_flags_off + xin_flag_bytes, 256);
  h->xin_ctx_off = align_up(sizeof(float) * (size_t)h->s.max_batch * h->s.dim, 256);
  h->xin_slot_bytes = h->xin_ctx_off + align_up(sizeof(uint64_t) * (size_t)h->s.max_batch, 256);
  const size_t total = h->xin_off + kPipeSlots * h->xin_slot_bytes;
  QRE_CUDA(cudaMalloc((void**)&h->xchg, total));
  QRE_CUDA(cudaMemset(h->xchg, 0, total));
  QRE_CUDA(cudaMalloc((void**)&h->xchg_ticket, 2 * sizeof(unsigned)));
  QRE_CUDA(cudaMemset(h->xchg_ticket, 0, 2 * sizeof(unsigned)));
  h->xin_ticket = h->xchg_ticket + 1;
  h->xin_epoch = 0;
  cudaIpcMemHandle_t ipc;
  QRE_CUDA(cudaIpcGetMemHandle(&ipc, h->xchg));
  memcpy(handle_out, &ipc, sizeof(ipc));
  h->xchg_world = world;
  h->xchg_rank = rank;
  h->xchg_peer[rank] = h->xchg;
  h->xchg_pub_epoch = h->xchg_merge_epoch = 0;
  return QRE_OK;
}

int qre_xchg_connect(qre_cache_t* h, int peer, const uint8_t* handle) {
  if (h == nullptr || handle == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (h->xchg == nullptr) return fail(QRE_ERR_INVALID, "call qre_xchg_init first");
  if (peer < 0 || peer >= h->xchg_world || peer == h->xchg_rank) return fail(QRE_ERR_INVALID, "bad peer %d", peer);
  DeviceGuard guard(h->s.device);
  cudaIpcMemHandle_t ipc;
  memcpy(&ipc, handle, sizeof(ipc));
  void* p = nullptr;
  QRE_CUDA(cudaIpcOpenMemHandle(&p, ipc, cudaIpcMemLazyEnablePeerAccess));
  h->xchg_peer[peer] = (uint8_t*)p;
  h->xchg_open[peer] = true;
  return QRE_OK;
}

static int publish_impl(qre_cache_t* h, int m, const qre_candidate_t* cand, void* stream, int per);

int qre_publish_dev(qre_cache_t* h, int m, const qre_candidate_t* cand, void* stream) {
  return publish_impl(h, m, cand, stream, 0);
}

// per > 0: peer p receives only the lists of queries [p * per, (p + 1) * per)
static int publish_impl(qre_cache_t* h, int m, const qre_candidate_t* cand, void* stream, int per) {
  if (h == nullptr || cand == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (h->xchg == nullptr) return fail(QRE_ERR_INVALID, "call qre_xchg_init / qre_xchg_connect first");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  for (int pr = 0; pr < h->xchg_world; ++pr)
    if (h->xchg_peer[pr] == nullptr) return fail(QRE_ERR_INVALID, "peer %d is not connected", pr);
  // One published batch may be outstanding per handle: publish(e) must be followed by merge(e)
  // before publish(e+1).  A peer's publish(e+2) reuses slot e&1 and is ordered (through its own
  // merge(e+1), which waits for my publish(e+1)) only after my merge(e) HAS BEEN ENQUEUED when the
  // calls alternate on one stream; a deeper host-side queue would let it overwrite lists my
  // merge(e) is still reading.
  if (h->xchg_pub_epoch != h->xchg_merge_epoch)
    return fail(QRE_ERR_INVALID, "batch %u is published but not merged: call qre_merge_route_xchg_dev before "
                "publishing again (publish and merge must alternate, on one stream)", h->xchg_pub_epoch);
  DeviceGuard guard(h->s.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : h->s.stream;
  const uint32_t epoch = ++h->xchg_pub_epoch;
  const int slot = (int)(epoch & 1u);
  qre_candidate_t* dst[16];
  uint32_t* flag[16];
  for (int pr = 0; pr < h->xchg_world; ++pr) {
    const size_t idx = (size_t)slot * h->xchg_world + h->xchg_rank;
    dst[pr] = reinterpret_cast<qre_candidate_t*>(h->xchg_peer[pr]) + idx * xchg_list_records(h);
    flag[pr] = reinterpret_cast<uint32_t*>(h->xchg_peer[pr] + h->xchg_flags_off) + idx * 32;
  }
  h->launches += launch_publish(m, h->xchg_world, cand, dst, flag, epoch, h->xchg_ticket, st, per);
  QRE_CUDA(cudaGetLastError());
  return QRE_OK;
}

int qre_scan_publish_dev(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx,
                         const qre_params_t* params, int32_t* overflow_count, void* stream) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (h->xchg == nullptr) return fail(QRE_ERR_INVALID, "call qre_xchg_init / qre_xchg_connect first");
  if (params == nullptr) return fail(QRE_ERR_INVALID, "params is NULL");
  qre_candidate_t* cand = h->s.batch[params->pipeline_slot & 1].cand;
  int rc = qre_scan_dev(h, m, queries, q_ctx, params, h->xchg_rank, cand, overflow_count, stream);
  if (rc != QRE_OK) return rc;
  return qre_publish_dev(h, m, cand, stream);
}

static int merge_xchg_impl(qre_cache_t* h, int m, const int32_t* tok, const uint8_t* heur_dev,
                           const double* heur_conf, const qre_params_t* params, qre_result_t* out_dev,
                           void* stream, int q0, int q1);

int qre_merge_route_xchg_dev(qre_cache_t* h, int m, const int32_t* tok, const uint8_t* heur_dev,
                             const double* heur_conf, const qre_params_t* params, qre_result_t* out_dev,
                             void* stream) {
  return merge_xchg_impl(h, m, tok, heur_dev, heur_conf, params, out_dev, stream, -1, -1);
}

static int merge_xchg_impl(qre_cache_t* h, int m, const int32_t* tok, const uint8_t* heur_dev,
                           const double* heur_conf, const qre_params_t* params, qre_result_t* out_dev,
                           void* stream, int q0, int q1) {
  if (h == nullptr || out_dev == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (h->xchg == nullptr) return fail(QRE_ERR_INVALID, "exchange not initialised");
  if (h->xchg_merge_epoch >= h->xchg_pub_epoch) return fail(QRE_ERR_INVALID, "nothing published to merge");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  int rc = validate_params(h, params);
  if (rc != QRE_OK) return rc;
  if (params->do_route && (!tok || !heur_dev || !heur_conf))
    return fail(QRE_ERR_INVALID, "do_route needs tok, heur_dev and heur_conf");
  DeviceGuard guard(h->s.device);
  cudaStream_t st = stream ? (cudaStream_t)stream : h->s.stream;
  const uint32_t epoch = ++h->xchg_merge_epoch;
  const int slot = (int)(epoch & 1u);
  const qre_candidate_t* lists =
      reinterpret_cast<const qre_candidate_t*>(h->xchg) + (size_t)slot * h->xchg_world * xchg_list_records(h);
  const uint32_t* flags = reinterpret_cast<const uint32_t*>(h->xchg + h->xchg_flags_off) +
                          (size_t)slot * h->xchg_world * 32;
  h->s.slot = params->pipeline_slot & 1;
  h->launches += launch_merge_route(h->s, m, h->xchg_world, lists, tok, heur_dev, heur_conf, *params, *out_dev, st,
                                    (long long)xchg_list_records(h), flags, epoch, q0, q1);
  QRE_CUDA(cudaGetLastError());
  return QRE_OK;
}

// ---- the pipeline -----------------------------------------------------------------------------
// One batch = prep (aux stream) -> scan (main stream, nothing else) -> re-score, settle, exchange,
// merge + route (post stream).  `in_ready` (may be NULL) is an event after which the inputs are
// valid.  Batch i uses pipeline slot i % kPipeSlots; that slot's previous batch (i - 3) must have
// left the post stream before its buffers are overwritten: the aux stream waits for its ev_post.
// Three slots, not two, so that the chain "re-score + exchange + merge of batch i-1 -> prep of
// batch i+1 -> scan of batch i+1" is NOT on the critical path: prep(i+1) only waits for batch
// i-2, which finished a whole scan earlier (with two slots that chain had to fit under one
// 250 us scan at 8 GPUs and left ~13 us gaps between scans).  The caller's output arrays belong
// to the API slot (0 / 1): the routing kernel of a batch, which writes them early, additionally
// waits for the batch previously submitted on the same API slot.  With `serial` everything is
// enqueued on the main stream (small batches are latency-bound: cross-stream hand-offs would
// only add to it).
static int pipeline_enqueue(qre_cache_t* h, int slot, int m, const float* d_q, const uint64_t* d_ctx,
                            const int32_t* d_tok, const uint8_t* d_hd, const double* d_hc, const qre_params_t& p_in,
                            const qre_result_t& out_dev, cudaEvent_t in_ready, bool serial, bool scatter_inputs) {
  DeviceState& s = h->s;
  qre_params_t p = p_in;
  const int api_slot = slot;
  const int prev_ps = h->api_ps[api_slot];                 // batch whose outputs this one overwrites
  slot = (int)(h->pipe_seq++ % kPipeSlots);                // from here on: the pipeline slot
  h->api_ps[api_slot] = slot;
  p.pipeline_slot = slot;
  s.slot = slot;
  h->last_slot = slot;
  h->last_m = m;
  DeviceState::Batch& B = s.cur();
  cudaStream_t aux = serial ? s.stream : h->aux, post = serial ? s.stream : h->post;
  cudaStream_t scan_st = (!serial && h->alt_scan && (h->pipe_seq & 1)) ? h->scan2 : s.stream;
  const bool sharded = h->xchg != nullptr && h->xchg_world > 1;
  // sliced results: this rank merges and routes queries [q0, q1) only
  const bool sliced = sharded && p.slice_results != 0;
  const int per = sliced ? (m + h->xchg_world - 1) / h->xchg_world : 0;
  const int q0 = sliced ? std::min(m, h->xchg_rank * per) : 0;
  const int q1 = sliced ? std::min(m, (h->xchg_rank + 1) * per) : m;
  if (in_ready) QRE_CUDA(cudaStreamWaitEvent(aux, in_ready, 0));
  if (h->post_used[slot]) QRE_CUDA(cudaStreamWaitEvent(aux, h->ev_post[slot], 0));
  if (scatter_inputs && sliced) {
    // only rows [q0, q1) of d_q / d_ctx were uploaded: store them into every rank's input buffer
    // (NVLink) and prepare the batch from this rank's copy once all slices have arrived
    for (int pr = 0; pr < h->xchg_world; ++pr)
      if (h->xchg_peer[pr] == nullptr) return fail(QRE_ERR_INVALID, "peer %d is not connected", pr);
    const uint32_t epoch = ++h->xin_epoch;
    const size_t xs = (size_t)slot;                        // same on every rank: they submit in lockstep
    float* dq[16];
    uint64_t* dc[16];
    uint32_t* fl[16];
    for (int pr = 0; pr < h->xchg_world; ++pr) {
      uint8_t* base = h->xchg_peer[pr] + h->xin_off + xs * h->xin_slot_bytes;
      dq[pr] = reinterpret_cast<float*>(base);
      dc[pr] = reinterpret_cast<uint64_t*>(base + h->xin_ctx_off);
      fl[pr] = reinterpret_cast<uint32_t*>(h->xchg_peer[pr] + h->xin_flags_off) + (xs * h->xchg_world + h->xchg_rank) * 32;
    }
    h->launches += launch_scatter_inputs(h->xchg_world, q0, q1, s.dim, d_q, d_ctx, dq, dc, fl, epoch, h->xin_ticket,
                                         aux);
    QRE_CUDA(cudaGetLastError());
    const uint8_t* mine = h->xchg + h->xin_off + xs * h->xin_slot_bytes;
    h->launches += launch_prep_queries(
        s, m, reinterpret_cast<const float*>(mine), reinterpret_cast<const uint64_t*>(mine + h->xin_ctx_off), aux,
        reinterpret_cast<const uint32_t*>(h->xchg + h->xin_flags_off) + xs * h->xchg_world * 32, h->xchg_world, epoch);
  } else {
    h->launches += launch_prep_queries(s, m, d_q, d_ctx, aux);
  }
  QRE_CUDA(cudaGetLastError());
  qre_params_t p_merge = p;
  if (serial && p.do_route) {
    // small batch: the chain stays on the main stream (launched as a programmatic-dependent chain),
    // only the routing epilogue forks to the aux stream and runs under the scan
    QRE_CUDA(cudaEventRecord(h->ev_prep[slot], s.stream));
    QRE_CUDA(cudaStreamWaitEvent(h->aux, h->ev_prep[slot], 0));
    if (prev_ps >= 0 && prev_ps != slot && h->post_used[prev_ps])
      QRE_CUDA(cudaStreamWaitEvent(h->aux, h->ev_post[prev_ps], 0));
    h->launches += launch_route_only(s, m, d_tok, d_hd, d_hc, p, out_dev, h->aux, q0, q1);
    QRE_CUDA(cudaGetLastError());
    QRE_CUDA(cudaEventRecord(h->ev_route[slot], h->aux));
    p_merge.do_route = 0;
  }
  s.pdl = serial;
  if (!serial) {
    QRE_CUDA(cudaEventRecord(h->ev_prep[slot], aux));
    QRE_CUDA(cudaStreamWaitEvent(scan_st, h->ev_prep[slot], 0));
    if (p.do_route) {
      // routing needs the prepared queries only: it runs on the aux stream beside the scan, and
      // the post stream's merge then fills in the lookup half of the results
      if (prev_ps >= 0 && prev_ps != slot && h->post_used[prev_ps])
        QRE_CUDA(cudaStreamWaitEvent(aux, h->ev_post[prev_ps], 0));
      h->launches += launch_route_only(s, m, d_tok, d_hd, d_hc, p, out_dev, aux, q0, q1);
      QRE_CUDA(cudaGetLastError());
      QRE_CUDA(cudaEventRecord(h->ev_route[slot], aux));
      p_merge.do_route = 0;
    }
  }
  int rc = tensor_scan_enqueue(h, m, p, scan_st);
  if (rc != QRE_OK) { s.pdl = false; return rc; }
  if (!serial) {
    QRE_CUDA(cudaEventRecord(h->ev_scan[slot], scan_st));
    QRE_CUDA(cudaStreamWaitEvent(post, h->ev_scan[slot], 0));
  }
  rc = rescore_enqueue(h, m, p, sharded ? h->xchg_rank : 0, B.cand, post);
  if (rc != QRE_OK) { s.pdl = false; return rc; }
  if (sharded) {
    s.pdl = false;                           // the exchange kernels are plain launches
    rc = publish_impl(h, m, B.cand, post, per);
    if (rc != QRE_OK) return rc;
    rc = merge_xchg_impl(h, m, d_tok, d_hd, d_hc, &p_merge, const_cast<qre_result_t*>(&out_dev), post, q0, q1);
    if (rc != QRE_OK) return rc;
  } else {
    h->launches += launch_merge_route(s, m, 1, B.cand, d_tok, d_hd, d_hc, p_merge, out_dev, post);
    s.pdl = false;
    QRE_CUDA(cudaGetLastError());
  }
  s.pdl = false;
  if (p.do_route) QRE_CUDA(cudaStreamWaitEvent(post, h->ev_route[slot], 0));   // batch done = both halves
  QRE_CUDA(cudaEventRecord(h->ev_post[slot], post));
  h->post_used[slot] = true;
  return QRE_OK;
}

int qre_submit_dev(qre_cache_t* h, int slot, int m, const float* queries, const uint64_t* q_ctx, const int32_t* tok,
                   const uint8_t* heur_dev, const double* heur_conf, const qre_params_t* params,
                   qre_result_t* out_dev, void* in_stream) {
  if (h == nullptr || out_dev == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (slot < 0 || slot >= kApiSlots) return fail(QRE_ERR_INVALID, "slot must be 0, 1 or 2");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  if (!queries || !q_ctx) return fail(QRE_ERR_INVALID, "NULL argument");
  int rc = validate_params(h, params);
  if (rc != QRE_OK) return rc;
  if (params->do_route && (!tok || !heur_dev || !heur_conf))
    return fail(QRE_ERR_INVALID, "do_route needs tok, heur_dev and heur_conf");
  DeviceGuard guard(h->s.device);
  cudaEvent_t ready = nullptr;
  if (in_stream != nullptr) {
    QRE_CUDA(cudaEventRecord(h->ev_in[slot], (cudaStream_t)in_stream));
    ready = h->ev_in[slot];
  }
  return pipeline_enqueue(h, slot, m, queries, q_ctx, tok, heur_dev, heur_conf, *params, *out_dev, ready, false,
                          false);
}

int qre_wait_dev(qre_cache_t* h, int slot, void* stream) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (slot < 0 || slot >= kApiSlots) return fail(QRE_ERR_INVALID, "slot must be 0, 1 or 2");
  const int ps = h->api_ps[slot];
  if (ps < 0 || !h->post_used[ps]) return QRE_OK;
  DeviceGuard guard(h->s.device);
  if (stream != nullptr) QRE_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, h->ev_post[ps], 0));
  else QRE_CUDA(cudaEventSynchronize(h->ev_post[ps]));
  return QRE_OK;
}

int qre_lookup_route_begin(qre_cache_t* h, int slot, int m, const float* queries, const uint64_t* q_ctx,
                           const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                           const qre_params_t* params) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  if (slot < 0 || slot >= kApiSlots) return fail(QRE_ERR_INVALID, "slot must be 0, 1 or 2");
  if (m < 1 || m > h->s.max_batch) return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d]", m, h->s.max_batch);
  if (!queries || !q_ctx) return fail(QRE_ERR_INVALID, "NULL argument");
  int rc = validate_params(h, params);
  if (rc != QRE_OK) return rc;
  if (params->do_route && (!tok || !heur_dev || !heur_conf))
    return fail(QRE_ERR_INVALID, "do_route needs tok, heur_dev and heur_conf");
  qre_cache_t::HostSlot& S = h->hs[slot];
  if (S.pending) return fail(QRE_ERR_INVALID, "slot %d still has a batch in flight: call qre_lookup_route_end", slot);
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  S.m = m;
  S.p = *params;
  // ---- inputs ----
  // Large batches: small arrays through pinned staging, the query matrix straight from the caller's
  // buffer when that is already page-locked, one H2D copy on the copy stream (it runs under the
  // previous batch's scan) and one D2H copy of the result block on the other copy stream.
  // Small batches (one query block, ~120 us of scan) are latency-bound: no copies at all — the
  // kernels read the inputs from, and write the results to, page-locked host memory directly.
  const bool zero_copy = m <= kBM;
  const bool direct = is_pinned_host(queries);
  // Row-sharded cache with sliced results: this rank returns queries [q0, q1) and (large batches)
  // uploads only that slice of every input array; the ranks exchange the query slices over NVLink.
  const bool sharded = h->xchg != nullptr && h->xchg_world > 1;
  const bool sliced = sharded && S.p.slice_results != 0;
  const int per = sliced ? (m + h->xchg_world - 1) / h->xchg_world : 0;
  S.q0 = sliced ? std::min(m, h->xchg_rank * per) : 0;
  S.q1 = sliced ? std::min(m, (h->xchg_rank + 1) * per) : m;
  const bool slice_up = sliced && !zero_copy;
  const size_t u0 = slice_up ? (size_t)S.q0 : 0, un = slice_up ? (size_t)(S.q1 - S.q0) : (size_t)m;
  uint8_t* in_base = zero_copy ? S.h_in : S.d_in;      // pinned host memory is device-accessible (UVA)
  size_t o = 0;
  struct Piece { size_t off, bytes; const void* direct_src; };
  Piece pieces[5];
  int n_pieces = 0;
  // place one input array at the next offset; stage (and later upload) rows [u0, u0 + un) of it
  auto put = [&](const void* src, size_t elem, bool copy) -> uint8_t* {
    uint8_t* d = in_base + o;
    if (src != nullptr) {
      const size_t off = o + u0 * elem, bytes = un * elem;
      if (copy && bytes) memcpy(S.h_in + off, static_cast<const uint8_t*>(src) + u0 * elem, bytes);
      pieces[n_pieces++] = Piece{off, bytes, copy ? nullptr : static_cast<const uint8_t*>(src) + u0 * elem};
    }
    o = align_up(o + (size_t)m * elem, 256);
    return d;
  };
  const float* d_q = (const float*)put(queries, sizeof(float) * (size_t)s.dim, !direct);
  if (zero_copy && direct) d_q = queries;
  const uint64_t* d_ctx = (const uint64_t*)put(q_ctx, sizeof(uint64_t), true);
  S.d_tok = (const int32_t*)put(tok, sizeof(int32_t), true);
  S.d_hd = put(heur_dev, 1, true);
  S.d_hc = (const double*)put(heur_conf, sizeof(double), true);
  if (!zero_copy) {
    if (!slice_up && !direct) {
      QRE_CUDA(cudaMemcpyAsync(S.d_in, S.h_in, o, cudaMemcpyHostToDevice, h->h2d));     // one copy of everything
    } else if (!slice_up) {
      // page-locked query matrix straight from the caller; the small arrays in one staged copy
      QRE_CUDA(cudaMemcpyAsync(S.d_in, queries, pieces[0].bytes, cudaMemcpyHostToDevice, h->h2d));
      const size_t small_off = align_up(pieces[0].bytes, 256);
      if (o > small_off)
        QRE_CUDA(cudaMemcpyAsync(S.d_in + small_off, S.h_in + small_off, o - small_off, cudaMemcpyHostToDevice,
                                 h->h2d));
    } else {
      for (int i = 0; i < n_pieces; ++i) {
        if (pieces[i].bytes == 0) continue;
        const void* src = pieces[i].direct_src ? pieces[i].direct_src : (const void*)(S.h_in + pieces[i].off);
        QRE_CUDA(cudaMemcpyAsync(S.d_in + pieces[i].off, src, pieces[i].bytes, cudaMemcpyHostToDevice, h->h2d));
      }
    }
    QRE_CUDA(cudaEventRecord(S.e_in, h->h2d));
  }
  const int K = S.p.do_route ? s.n_centroids : 0;
  const int D = S.p.do_route ? S.p.n_devices : 1;
  S.L.plan(m, K, D);
  uint8_t* res_base = zero_copy ? S.h_res : S.d_res;
  const qre_result_t dres = S.L.view(res_base);
  rc = pipeline_enqueue(h, slot, m, d_q, d_ctx, S.d_tok, S.d_hd, S.d_hc, S.p, dres, zero_copy ? nullptr : S.e_in,
                        zero_copy, slice_up);
  if (rc != QRE_OK) return rc;
  if (zero_copy) {
    QRE_CUDA(cudaEventRecord(S.e_done, s.stream));
  } else {
    QRE_CUDA(cudaStreamWaitEvent(h->d2h, h->ev_post[h->api_ps[slot]], 0));
    QRE_CUDA(cudaMemcpyAsync(S.h_res, S.d_res, S.L.total, cudaMemcpyDeviceToHost, h->d2h));
    QRE_CUDA(cudaEventRecord(S.e_done, h->d2h));
  }
  S.pending = true;
  return QRE_OK;
}

int qre_lookup_route_end(qre_cache_t* h, int slot, qre_result_t* out) {
  if (h == nullptr || out == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (slot < 0 || slot >= kApiSlots) return fail(QRE_ERR_INVALID, "slot must be 0, 1 or 2");
  qre_cache_t::HostSlot& S = h->hs[slot];
  if (!S.pending) return fail(QRE_ERR_INVALID, "slot %d has no batch in flight", slot);
  DeviceGuard guard(h->s.device);
  QRE_CUDA(cudaEventSynchronize(S.e_done));
  S.pending = false;
  // every overflowed query has been settled on the device (the settle pass covers a whole batch),
  // so the asynchronous path is exact on its own
  const int K = S.p.do_route ? h->s.n_centroids : 0;
  const int D = S.p.do_route ? S.p.n_devices : 1;
  scatter_results(S, K, D, out);
  return QRE_OK;
}

int qre_lookup_route(qre_cache_t* h, int m, const float* queries, const uint64_t* q_ctx, const int32_t* tok,
                     const uint8_t* heur_dev, const double* heur_conf, const qre_params_t* params,
                     qre_result_t* out) {
  if (out == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (h != nullptr && (h->hs[0].pending || h->hs[1].pending || h->hs[2].pending))
    return fail(QRE_ERR_INVALID, "asynchronous batches in flight: finish them with qre_lookup_route_end first");
  int rc = qre_lookup_route_begin(h, 0, m, queries, q_ctx, tok, heur_dev, heur_conf, params);
  if (rc != QRE_OK) return rc;
  return qre_lookup_route_end(h, 0, out);
}

// ---- one process, several GPUs ---------------------------------------------------------------
// The row-sharded lookup for a host that drives all GPUs of the box from ONE process (the
// drop-in QueryCache with config["gpus"]): every handle scans and re-scores its rows for the
// whole batch on its own GPU and stream, the per-shard candidate lists are copied to the first
// handle's GPU (peer copies over NVLink), and the merge + routing kernel runs there over all of
// them — the same kernel that merges the ranks' lists in the multi-process path.
int qre_group_lookup_route(qre_cache_t* const* hs, int n_handles, int m, const float* queries,
                           const uint64_t* q_ctx, const int32_t* tok, const uint8_t* heur_dev,
                           const double* heur_conf, const qre_params_t* params, qre_result_t* out) {
  if (hs == nullptr || out == nullptr || queries == nullptr || q_ctx == nullptr)
    return fail(QRE_ERR_INVALID, "NULL argument");
  if (n_handles < 1 || n_handles * QRE_TOPK > 128) return fail(QRE_ERR_INVALID, "n_handles out of range [1, 16]");
  for (int i = 0; i < n_handles; ++i) {
    if (hs[i] == nullptr) return fail(QRE_ERR_INVALID, "handle %d is NULL", i);
    if (hs[i]->s.dim != hs[0]->s.dim) return fail(QRE_ERR_INVALID, "handle %d has another dimension", i);
    if (m < 1 || m > hs[i]->s.max_batch)
      return fail(QRE_ERR_CAPACITY, "m=%d outside [1, max_batch=%d] of handle %d", m, hs[i]->s.max_batch, i);
    if (hs[i]->hs[0].pending || hs[i]->hs[1].pending || hs[i]->hs[2].pending)
      return fail(QRE_ERR_INVALID, "handle %d has asynchronous batches in flight", i);
  }
  qre_cache_t* h0 = hs[0];
  int rc = validate_params(h0, params);
  if (rc != QRE_OK) return rc;
  if (params->do_route && (!tok || !heur_dev || !heur_conf))
    return fail(QRE_ERR_INVALID, "do_route needs tok, heur_dev and heur_conf");
  const int dim = h0->s.dim;
  const size_t list_bytes = sizeof(qre_candidate_t) * (size_t)m * QRE_TOPK;
  const size_t stride = (size_t)h0->s.max_batch * QRE_TOPK;
  {
    DeviceGuard g0(h0->s.device);
    if (h0->d_gather == nullptr) {
      QRE_CUDA(cudaMalloc((void**)&h0->d_gather, sizeof(qre_candidate_t) * 16 * stride));
      QRE_CUDA(cudaEventCreateWithFlags(&h0->ev_group, cudaEventDisableTiming));
    }
  }
  qre_params_t p = *params;
  p.pipeline_slot = 0;
  p.slice_results = 0;
  const int32_t* d_tok0 = nullptr;
  const uint8_t* d_hd0 = nullptr;
  const double* d_hc0 = nullptr;
  // ---- every shard: inputs in, prep + scan + exact re-score on its own stream, list to GPU 0 ----
  for (int i = 0; i < n_handles; ++i) {
    qre_cache_t* h = hs[i];
    DeviceGuard guard(h->s.device);
    qre_cache_t::HostSlot& S = h->hs[0];
    size_t o = 0;
    auto put = [&](const void* src, size_t bytes) -> uint8_t* {
      uint8_t* d = S.d_in + o;
      if (src) memcpy(S.h_in + o, src, bytes);
      o = align_up(o + bytes, 256);
      return d;
    };
    const float* d_q = (const float*)put(queries, sizeof(float) * (size_t)m * dim);
    const uint64_t* d_ctx = (const uint64_t*)put(q_ctx, sizeof(uint64_t) * (size_t)m);
    if (i == 0) {                                   // routing inputs are only read where the merge runs
      d_tok0 = (const int32_t*)put(tok, sizeof(int32_t) * (size_t)m);
      d_hd0 = put(heur_dev, (size_t)m);
      d_hc0 = (const double*)put(heur_conf, sizeof(double) * (size_t)m);
    }
    cudaStream_t st = h->s.stream;
    QRE_CUDA(cudaMemcpyAsync(S.d_in, S.h_in, o, cudaMemcpyHostToDevice, st));
    h->s.slot = 0;
    rc = scan_enqueue(h, m, d_q, d_ctx, p, i, h->s.batch[0].cand, st);
    if (rc != QRE_OK) return rc;
    if (i == 0) {
      QRE_CUDA(cudaMemcpyAsync(h0->d_gather, h->s.batch[0].cand, list_bytes, cudaMemcpyDeviceToDevice, st));
    } else {
      QRE_CUDA(cudaMemcpyPeerAsync(h0->d_gather + (size_t)i * stride, h0->s.device, h->s.batch[0].cand,
                                   h->s.device, list_bytes, st));
      QRE_CUDA(cudaEventRecord(h->ev_scan[0], st));
    }
  }
  // ---- GPU 0: merge the n lists per query, route, results out ----
  DeviceGuard g0(h0->s.device);
  DeviceState& s0 = h0->s;
  qre_cache_t::HostSlot& S0 = h0->hs[0];
  for (int i = 1; i < n_handles; ++i) QRE_CUDA(cudaStreamWaitEvent(s0.stream, hs[i]->ev_scan[0], 0));
  S0.m = m; S0.p = p; S0.q0 = 0; S0.q1 = m;
  const int K = p.do_route ? s0.n_centroids : 0;
  const int D = p.do_route ? p.n_devices : 1;
  S0.L.plan(m, K, D);
  s0.slot = 0;
  h0->launches += launch_merge_route(s0, m, n_handles, h0->d_gather, d_tok0, d_hd0, d_hc0, p, S0.L.view(S0.d_res),
                                     s0.stream, (long long)stride);
  QRE_CUDA(cudaGetLastError());
  QRE_CUDA(cudaMemcpyAsync(S0.h_res, S0.d_res, S0.L.total, cudaMemcpyDeviceToHost, s0.stream));
  QRE_CUDA(cudaStreamSynchronize(s0.stream));
  scatter_results(S0, K, D, out);
  for (int i = 1; i < n_handles; ++i) {              // nothing of this batch may be left in flight
    DeviceGuard guard(hs[i]->s.device);
    QRE_CUDA(cudaStreamSynchronize(hs[i]->s.stream));
  }
  return QRE_OK;
}

// ---- binary snapshot -----------------------------------------------------
// Layout: header | slot_live[hwm] | RowMeta[hwm] | lru[hwm] | pred_dev[hwm] | pred_conf[hwm] |
// rows_f32[hwm][dim] (unpadded).  Dead slots are stored too (their contents are ignored on load).
namespace {
struct SnapHeader {
  char magic[8];
  int32_t version, dim;
  int64_t hwm, live;
};
}  // namespace

int qre_cache_snapshot_save(qre_cache_t* h, const char* path) {
  if (h == nullptr || path == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  FILE* f = fopen(path, "wb");
  if (f == nullptr) return fail(QRE_ERR_INVALID, "cannot open %s for writing", path);
  SnapHeader hd;
  memcpy(hd.magic, "QRESNAP1", 8);
  hd.version = 1; hd.dim = s.dim; hd.hwm = h->hwm; hd.live = h->live;
  const size_t n = (size_t)h->hwm;
  bool ok = fwrite(&hd, sizeof(hd), 1, f) == 1;
  if (n) ok = ok && fwrite(h->slot_live.data(), 1, n, f) == n;
  int rc = QRE_OK;
  auto dump = [&](const void* dev, size_t bytes) -> int {
    std::vector<uint8_t> buf(bytes);
    QRE_CUDA(cudaMemcpyAsync(buf.data(), dev, bytes, cudaMemcpyDeviceToHost, s.stream));
    QRE_CUDA(cudaStreamSynchronize(s.stream));
    ok = ok && (bytes == 0 || fwrite(buf.data(), 1, bytes, f) == bytes);
    return QRE_OK;
  };
  if (n) {
    if (rc == QRE_OK) rc = dump(s.meta, sizeof(RowMeta) * n);
    if (rc == QRE_OK) rc = dump(s.lru, sizeof(uint64_t) * n);
    if (rc == QRE_OK) rc = dump(s.pred_dev, n);
    if (rc == QRE_OK) rc = dump(s.pred_conf, sizeof(double) * n);
    const size_t chunk = std::max<size_t>(1, (32u << 20) / (sizeof(float) * s.dim));
    std::vector<float> buf(chunk * s.dim);
    for (size_t r0 = 0; r0 < n && rc == QRE_OK && ok; r0 += chunk) {
      const size_t c = std::min(chunk, n - r0);
      cudaError_t e = cudaMemcpy2DAsync(buf.data(), sizeof(float) * s.dim, s.rows_f32 + r0 * s.dim_pad,
                                        sizeof(float) * s.dim_pad, sizeof(float) * s.dim, c,
                                        cudaMemcpyDeviceToHost, s.stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
      if (e != cudaSuccess) { rc = fail(QRE_ERR_CUDA, "snapshot copy failed: %s", cudaGetErrorString(e)); break; }
      ok = ok && fwrite(buf.data(), sizeof(float) * s.dim, c, f) == c;
    }
  }
  ok = (fclose(f) == 0) && ok;
  if (rc != QRE_OK) return rc;
  if (!ok) return fail(QRE_ERR_INVALID, "short write to %s", path);
  return QRE_OK;
}

int qre_cache_snapshot_load(qre_cache_t* h, const char* path) {
  if (h == nullptr || path == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (h->hwm != 0) return fail(QRE_ERR_INVALID, "snapshot_load needs an empty cache (call qre_cache_clear)");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  FILE* f = fopen(path, "rb");
  if (f == nullptr) return fail(QRE_ERR_INVALID, "cannot open %s", path);
  SnapHeader hd;
  int rc = QRE_OK;
  if (fread(&hd, sizeof(hd), 1, f) != 1 || memcmp(hd.magic, "QRESNAP1", 8) != 0 || hd.version != 1)
    rc = fail(QRE_ERR_INVALID, "%s is not a qre snapshot", path);
  else if (hd.dim != s.dim) rc = fail(QRE_ERR_INVALID, "snapshot dim %d != cache dim %d", hd.dim, s.dim);
  else if (hd.hwm < 0 || hd.hwm > s.capacity)
    rc = fail(QRE_ERR_CAPACITY, "snapshot holds %lld slots, capacity is %lld", (long long)hd.hwm, s.capacity);
  if (rc != QRE_OK) { fclose(f); return rc; }
  const size_t n = (size_t)hd.hwm;
  std::vector<uint8_t> live(n);
  std::vector<RowMeta> meta(n);
  std::vector<uint64_t> lru(n);
  std::vector<uint8_t> pdev(n);
  std::vector<double> pconf(n);
  bool ok = true;
  if (n) {
    ok = fread(live.data(), 1, n, f) == n && fread(meta.data(), sizeof(RowMeta), n, f) == n &&
         fread(lru.data(), sizeof(uint64_t), n, f) == n && fread(pdev.data(), 1, n, f) == n &&
         fread(pconf.data(), sizeof(double), n, f) == n;
  }
  const size_t chunk = std::max<size_t>(1, (32u << 20) / (sizeof(float) * s.dim));
  std::vector<float> rows(chunk * s.dim), rows_c(chunk * s.dim);
  std::vector<int64_t> slot;
  std::vector<uint64_t> ctx_c, lru_c;
  std::vector<int64_t> ts_c;
  std::vector<uint8_t> pdev_c;
  std::vector<double> pconf_c;
  long long n_live = 0;
  for (size_t r0 = 0; r0 < n && ok && rc == QRE_OK; r0 += chunk) {
    const size_t c = std::min(chunk, n - r0);
    ok = fread(rows.data(), sizeof(float) * s.dim, c, f) == c;
    if (!ok) break;
    slot.clear(); ctx_c.clear(); lru_c.clear(); ts_c.clear(); pdev_c.clear(); pconf_c.clear();
    size_t k = 0;
    for (size_t i = 0; i < c; ++i) {
      const size_t g = r0 + i;
      if (!live[g]) continue;
      memcpy(rows_c.data() + k * s.dim, rows.data() + i * s.dim, sizeof(float) * s.dim);
      slot.push_back((int64_t)g); ctx_c.push_back(meta[g].ctx); ts_c.push_back(meta[g].ts_us);
      lru_c.push_back(lru[g]); pdev_c.push_back(pdev[g]); pconf_c.push_back(pconf[g]);
      ++k;
    }
    if (k)
      rc = staged_write(h, (int64_t)k, slot.data(), rows_c.data(), ctx_c.data(), ts_c.data(), lru_c.data(),
                        pdev_c.data(), pconf_c.data());
    n_live += (long long)k;
  }
  fclose(f);
  if (rc != QRE_OK) return rc;
  if (!ok) { qre_cache_clear(h); return fail(QRE_ERR_INVALID, "%s is truncated", path); }
  h->hwm = hd.hwm;
  h->live = n_live;
  h->free_slots.clear();
  for (size_t g = n; g-- > 0;) {
    h->slot_live[g] = live[g];
    if (!live[g]) h->free_slots.push_back((int64_t)g);
  }
  return QRE_OK;
}

int qre_profile(qre_cache_t* h, int enable) {
  if (h == nullptr) return fail(QRE_ERR_INVALID, "handle is NULL");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  if (enable && s.prof_ev == nullptr) {
    s.prof_ev = new (std::nothrow) cudaEvent_t[2 * kProfMax];
    if (s.prof_ev == nullptr) return fail(QRE_ERR_INVALID, "out of host memory");
    for (int i = 0; i < 2 * kProfMax; ++i) QRE_CUDA(cudaEventCreate(&s.prof_ev[i]));
  }
  s.prof_on = enable != 0;
  s.prof_used = 0;
  return QRE_OK;
}

int qre_profile_read(qre_cache_t* h, double* scan_ms_total, int* n_launches) {
  if (h == nullptr || scan_ms_total == nullptr || n_launches == nullptr)
    return fail(QRE_ERR_INVALID, "NULL argument");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  // Consecutive scans run on alternating streams and overlap at their edges (the next one's start
  // event fires while the previous kernel still holds the SMs), so a launch is charged from its own
  // start event or the previous launch's end event, whichever is later, to its own end event: the
  // sum is the time the scan kernels hold the GPU, not the sum of overlapping brackets.
  double total = 0.0;
  for (int i = 0; i < s.prof_used; ++i) {
    float ms = 0.f;
    QRE_CUDA(cudaEventSynchronize(s.prof_ev[2 * i + 1]));
    QRE_CUDA(cudaEventElapsedTime(&ms, s.prof_ev[2 * i], s.prof_ev[2 * i + 1]));
    if (i > 0) {
      float since_prev = 0.f;
      QRE_CUDA(cudaEventElapsedTime(&since_prev, s.prof_ev[2 * i - 1], s.prof_ev[2 * i + 1]));
      if (since_prev >= 0.f && since_prev < ms) ms = since_prev;
    }
    total += ms;
  }
  *scan_ms_total = total;
  *n_launches = s.prof_used;
  return QRE_OK;
}

int qre_debug_gemm(qre_cache_t* h, int m, int64_t row0, int n, float* out) {
  if (h == nullptr || out == nullptr) return fail(QRE_ERR_INVALID, "NULL argument");
  if (m < 1 || m != h->last_m) return fail(QRE_ERR_INVALID, "run a scan with the same m first");
  if (n < 1 || row0 < 0 || row0 + n > h->hwm) return fail(QRE_ERR_INVALID, "row range outside the cache");
  DeviceGuard guard(h->s.device);
  DeviceState& s = h->s;
  const size_t bytes = sizeof(float) * (size_t)m * n;
  if (bytes > h->dump_bytes) {
    if (h->d_dump) cudaFree(h->d_dump);
    h->d_dump = nullptr;
    h->dump_bytes = 0;
    QRE_CUDA(cudaMalloc((void**)&h->d_dump, bytes));
    h->dump_bytes = bytes;
  }
  QRE_CUDA(cudaMemsetAsync(h->d_dump, 0, bytes, s.stream));
  qre_params_t p;
  memset(&p, 0, sizeof(p));
  p.similarity_threshold = 2.0;   // floor above any cosine: the candidate path stays idle
  p.ttl_us = 1;
  s.slot = h->last_slot;
  {
    const int nl = launch_gemm_topk(s, m, h->hwm, p, h->d_dump, n, row0, n, s.stream);
    if (nl < 0) return fail(QRE_ERR_CAPACITY, "scan geometry does not fit the top-k scratch");
    h->launches += nl;
  }
  QRE_CUDA(cudaGetLastError());
  QRE_CUDA(cudaMemcpyAsync(out, h->d_dump, bytes, cudaMemcpyDeviceToHost, s.stream));
  QRE_CUDA(cudaStreamSynchronize(s.stream));
  return QRE_OK;
}

}  // extern "C"
