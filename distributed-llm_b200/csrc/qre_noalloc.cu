// qre_noalloc.cu — qre_exact_lookup_noalloc: the brute-force exact lookup for callers that
// must not allocate GPU memory.
//
// Who this is for.  Some hosts of this library are not free to call cudaMalloc when a lookup
// arrives: a serving process that handed all of HBM to its own pool at start-up and treats
// any later driver allocation as a budget violation; a request thread that cannot afford the
// device-wide synchronisation cudaMalloc/cudaFree imply while other streams are busy; a
// plugin loaded into a framework whose rules forbid allocations outside the framework's
// allocator.  The handle-based API (qre_cache_create ...) allocates its scratch up front with
// cudaMalloc; this entry point instead takes every input and output from the caller and keeps
// the one piece of scratch it needs — the per-CTA partial winners of the scan — in a
// file-scope `static __device__` array that is part of the loaded module.
//
// Cost and limits of that choice: the array (kScratchBytes = 197,120 bytes) is resident on
// every GPU that loads the library until the process exits; at most QRE_NOALLOC_MAX_QUERIES
// (8) queries per call, dim <= 1024; every call on a device shares the same array, so two
// calls that overlap on that GPU (different streams or threads) race — serialise them.
//
// It is a variant of exact_scan_kernel (qre_kernels.cu): same masks, same fp32 summation
// order and rounding sequence, hence bit-identical scores; the cross-CTA reduction is folded
// into the same launch (last CTA to finish reduces), which needs a ticket counter that must
// start at zero — it lives at the head of the scratch array and is cleared before every launch.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/qre_b200.h"

namespace {

constexpr int kMaxQ = QRE_NOALLOC_MAX_QUERIES;
constexpr int kMaxDim = QRE_NOALLOC_MAX_DIM;
constexpr int kMaxCtas = 1024;
constexpr int kWarps = 8;
constexpr int kMaxIt = kMaxDim / 128;
constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr float kNormEps = 1e-9f;

struct Best {               // one CTA's (or the final) winner for one query
  float sim;
  int valid;
  long long row;
  unsigned long long lru;
};

constexpr size_t kHeaderBytes = 256;    // ticket counter (must be zero at launch)
constexpr size_t kScratchBytes = kHeaderBytes + sizeof(Best) * kMaxQ * kMaxCtas + 256;
static_assert(kScratchBytes == 197120, "documented size");

// The scratch.  In host code this name is only a stand-in for the symbol: the device address
// is obtained once per device with cudaGetSymbolAddress (see scratch_for_device) and passed
// to the kernel as an ordinary pointer.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

__device__ __forceinline__ float fma4(float4 x, float4 y, float acc) {
  acc = fmaf(x.x, y.x, acc);
  acc = fmaf(x.y, y.y, acc);
  acc = fmaf(x.z, y.z, acc);
  acc = fmaf(x.w, y.w, acc);
  return acc;
}
__device__ __forceinline__ float wsum(float acc) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
  return acc;
}
__device__ __forceinline__ bool better(float s0, unsigned long long l0, float s1, unsigned long long l1) {
  return s0 > s1 || (s0 == s1 && l0 < l1);      // (sim desc, lru asc): src/cache.py:292
}

struct Args {
  long long n_rows;
  int dim, m;
  const float* rows;
  const unsigned long long* row_ctx;
  const long long* row_ts;
  const unsigned long long* row_lru;
  const float* queries;
  const unsigned long long* q_ctx;
  double thr;
  long long now_us, ttl_us;
  long long* hit_row;
  float* hit_sim;
  unsigned char* scratch;
};

__global__ void __launch_bounds__(kWarps * 32) exact_lookup_noalloc_kernel(const Args a) {
  extern __shared__ __align__(16) float qs[];         // [m, dim], read back as float4
  __shared__ float s_nq[kMaxQ];
  __shared__ unsigned long long s_ctx[kMaxQ];
  __shared__ Best s_best[kWarps][kMaxQ];
  __shared__ unsigned s_ticket;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned* counter = reinterpret_cast<unsigned*>(a.scratch);
  Best* partial = reinterpret_cast<Best*>(a.scratch + kHeaderBytes);

  for (int i = threadIdx.x; i < a.m * a.dim; i += blockDim.x) qs[i] = a.queries[i];
  __syncthreads();
  if (warp < a.m) {                                   // |q| with the library's summation order
    float acc = 0.f;
    for (int i = lane * 4; i < a.dim; i += 128) {
      const float4 x = *reinterpret_cast<const float4*>(qs + warp * a.dim + i);
      acc = fma4(x, x, acc);
    }
    acc = wsum(acc);
    if (lane == 0) {
      s_nq[warp] = __fsqrt_rn(acc);
      s_ctx[warp] = a.q_ctx[warp];
    }
  }
  __syncthreads();

  const long long r0 = a.n_rows * blockIdx.x / gridDim.x, r1 = a.n_rows * (blockIdx.x + 1) / gridDim.x;
  float bsim = -INFINITY;
  long long brow = -1;
  unsigned long long blru = ~0ull;
  for (long long row = r0 + warp; row < r1; row += kWarps) {
    const unsigned long long ctx = a.row_ctx[row];
    if (ctx == QRE_CTX_TOMBSTONE || (a.now_us - a.row_ts[row]) > a.ttl_us) continue;    // cache.py:275-276
    const bool mine = lane < a.m && s_ctx[lane] == ctx && s_nq[lane] >= kNormEps;        // cache.py:274,284
    if (!__any_sync(kFull, mine)) continue;
    float4 e[kMaxIt];
    const float* rp = a.rows + row * a.dim;
    float nacc = 0.f;
#pragma unroll
    for (int it = 0; it < kMaxIt; ++it) {
      const int i = lane * 4 + it * 128;
      e[it] = i < a.dim ? *reinterpret_cast<const float4*>(rp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < a.dim) nacc = fma4(e[it], e[it], nacc);
    }
    const float ne = __fsqrt_rn(wsum(nacc));
    if (ne < kNormEps) continue;                                                          // cache.py:289
    float mydot = 0.f;
    for (int j = 0; j < a.m; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int it = 0; it < kMaxIt; ++it) {
        const int i = lane * 4 + it * 128;
        if (i < a.dim) acc = fma4(*reinterpret_cast<const float4*>(qs + j * a.dim + i), e[it], acc);
      }
      acc = wsum(acc);
      if (lane == j) mydot = acc;
    }
    if (mine) {
      const float sim = __fdiv_rn(mydot, __fmul_rn(s_nq[lane], ne));                      // cache.py:291
      if (static_cast<double>(sim) >= a.thr && sim > 0.f) {                               // cache.py:292
        const unsigned long long l = a.row_lru[row];
        if (better(sim, l, bsim, blru)) { bsim = sim; blru = l; brow = row; }
      }
    }
  }
  if (lane < kMaxQ) s_best[warp][lane] = Best{bsim, brow >= 0 ? 1 : 0, brow, blru};
  __syncthreads();
  if (warp == 0 && lane < a.m) {
    Best b = s_best[0][lane];
    for (int w = 1; w < kWarps; ++w) {
      const Best c = s_best[w][lane];
      if (c.valid && (!b.valid || better(c.sim, c.lru, b.sim, b.lru))) b = c;
    }
    partial[static_cast<size_t>(lane) * gridDim.x + blockIdx.x] = b;
  }
  // last CTA to finish reduces all partials (ticket counter was cleared before the launch)
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(counter, 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence();
  if (warp < a.m) {
    Best b{-INFINITY, 0, -1, ~0ull};
    for (unsigned i = lane; i < gridDim.x; i += 32) {
      const Best* p = &partial[static_cast<size_t>(warp) * gridDim.x + i];
      Best c;                                         // written by other CTAs: read through L2
      c.sim = __ldcg(&p->sim);
      c.valid = __ldcg(&p->valid);
      c.row = __ldcg(&p->row);
      c.lru = __ldcg(&p->lru);
      if (c.valid && (!b.valid || better(c.sim, c.lru, b.sim, b.lru))) b = c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      Best c;
      c.sim = __shfl_xor_sync(kFull, b.sim, o);
      c.valid = __shfl_xor_sync(kFull, b.valid, o);
      c.row = __shfl_xor_sync(kFull, b.row, o);
      c.lru = __shfl_xor_sync(kFull, b.lru, o);
      if (c.valid && (!b.valid || better(c.sim, c.lru, b.sim, b.lru))) b = c;
    }
    if (lane == 0) {
      a.hit_row[warp] = b.valid ? b.row : -1;
      a.hit_sim[warp] = b.valid ? b.sim : 0.f;
    }
  }
  if (threadIdx.x == 0) *counter = 0u;               // leave the ticket clean as well
}

// Device address of the scratch array, resolved once per device with that device current.
constexpr int kMaxGpus = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxGpus] = {};
int g_sms[kMaxGpus] = {};

cudaError_t scratch_for_device(int device, unsigned char** out, int* sms) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_scratch_ptr[device] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) return e;
    g_sms[device] = n;
    g_scratch_ptr[device] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_ptr[device];
  *sms = g_sms[device];
  return cudaSuccess;
}

int report(char* err, size_t err_len, int code, const char* msg, const char* detail = "") {
  if (err != nullptr && err_len > 0) snprintf(err, err_len, "%s%s", msg, detail);
  return code;
}

}  // namespace

extern "C" QRE_API int qre_exact_lookup_noalloc(
    int device, int64_t n_rows, int dim, const float* rows, const uint64_t* row_ctx, const int64_t* row_ts_us,
    const uint64_t* row_lru, int m, const float* queries, const uint64_t* q_ctx, double similarity_threshold,
    int64_t now_us, int64_t ttl_us, int64_t* hit_row, float* hit_sim, void* stream, char* err, size_t err_len) {
  if (m < 1 || m > kMaxQ)
    return report(err, err_len, QRE_ERR_CAPACITY, "qre_exact_lookup_noalloc: m must be in [1, 8] queries per call");
  if (dim < 4 || dim > kMaxDim || dim % 4 != 0)
    return report(err, err_len, QRE_ERR_CAPACITY,
                  "qre_exact_lookup_noalloc: dim must be a multiple of 4 in [4, 1024]");
  if (n_rows < 0 || n_rows > QRE_NOALLOC_MAX_ROWS)
    return report(err, err_len, QRE_ERR_CAPACITY, "qre_exact_lookup_noalloc: n_rows must be in [0, 2^31-1]");
  if (!queries || !q_ctx || !hit_row || !hit_sim || (n_rows > 0 && (!rows || !row_ctx || !row_ts_us || !row_lru)))
    return report(err, err_len, QRE_ERR_INVALID, "qre_exact_lookup_noalloc: NULL array");
  if (n_rows > 0 && (reinterpret_cast<uintptr_t>(rows) & 15u) != 0)
    return report(err, err_len, QRE_ERR_INVALID,
                  "qre_exact_lookup_noalloc: rows must be 16-byte aligned (rows are read as float4)");
  if (device < 0 || device >= kMaxGpus)
    return report(err, err_len, QRE_ERR_INVALID, "qre_exact_lookup_noalloc: device index out of range");
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
  if (e != cudaSuccess) return report(err, err_len, QRE_ERR_CUDA, "cudaSetDevice: ", cudaGetErrorString(e));
  unsigned char* scratch = nullptr;
  int sms = 0;
  int rc = QRE_OK;
  e = scratch_for_device(device, &scratch, &sms);
  if (e != cudaSuccess) {
    rc = report(err, err_len, QRE_ERR_CUDA, "cudaGetSymbolAddress(scratch): ", cudaGetErrorString(e));
  } else {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Args a;
    a.n_rows = n_rows; a.dim = dim; a.m = m; a.rows = rows;
    a.row_ctx = reinterpret_cast<const unsigned long long*>(row_ctx);
    a.row_ts = reinterpret_cast<const long long*>(row_ts_us);
    a.row_lru = reinterpret_cast<const unsigned long long*>(row_lru);
    a.queries = queries; a.q_ctx = reinterpret_cast<const unsigned long long*>(q_ctx);
    a.thr = similarity_threshold; a.now_us = now_us; a.ttl_us = ttl_us;
    a.hit_row = reinterpret_cast<long long*>(hit_row); a.hit_sim = hit_sim; a.scratch = scratch;
    long long grid = 2LL * sms;
    if (grid > kMaxCtas) grid = kMaxCtas;
    if (grid > n_rows) grid = n_rows;
    if (grid < 1) grid = 1;
    // the shared array keeps whatever the previous call left: the ticket must start at zero
    e = cudaMemsetAsync(scratch, 0, kHeaderBytes, st);
    if (e == cudaSuccess) {
      exact_lookup_noalloc_kernel<<<static_cast<unsigned>(grid), kWarps * 32,
                                    static_cast<size_t>(m) * dim * sizeof(float), st>>>(a);
      e = cudaGetLastError();
    }
    if (e != cudaSuccess) rc = report(err, err_len, QRE_ERR_CUDA, "launch failed: ", cudaGetErrorString(e));
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return rc;
}
