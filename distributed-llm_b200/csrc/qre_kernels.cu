// qre_kernels.cu — everything around the tensor-core scan: cache-row maintenance, query
// preparation, exact fp32 re-scoring of candidates, the brute-force fallback scan, and the
// merge + routing epilogue.  All of these are HBM/latency-bound CUDA-core kernels.
//
// Reference arithmetic restated here (file:line relative to /root/reference/src/):
//   cache.py:283-292  sim = fl32(sdot32(q,e) / fl32(|q|*|e|)); hit iff double(sim) >= thr
//                     and sim > 0; first strictly-greatest in OrderedDict order wins.
//   query_router_engine.py:133-138 (centroid cosine), 98-99 (token), 180-213 (semantic),
//   382-414 (hybrid vote).
// Double-precision steps use __dmul_rn/__dadd_rn so the compiler cannot contract them into
// FMAs: the reference rounds after every operation.
#include <math.h>
#include <stdio.h>

#include "qre_common.cuh"
#include "qre_encoder.cuh"      // pdl_wait / pdl_launch_dependents / launch_chain

namespace qre {

constexpr int kWarpsPerBlock = 4;
constexpr unsigned kFull = 0xFFFFFFFFu;

// fp32 dot product by one warp: lane l accumulates elements [4l + 128 it, +4) with FMAs in
// index order, then a 5-step xor tree.  Every norm and every exact score in the library
// goes through this one summation order.
__device__ __forceinline__ float lane_fma4(float4 x, float4 y, float acc) {
  acc = fmaf(x.x, y.x, acc);
  acc = fmaf(x.y, y.y, acc);
  acc = fmaf(x.z, y.z, acc);
  acc = fmaf(x.w, y.w, acc);
  return acc;
}
__device__ __forceinline__ float warp_sum(float acc) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
  return acc;
}
__device__ __forceinline__ float warp_dot(const float* __restrict__ a, const float* __restrict__ b, int n,
                                          int lane) {
  float acc = 0.f;
  for (int i = lane * 4; i < n; i += 128)
    acc = lane_fma4(*reinterpret_cast<const float4*>(a + i), *reinterpret_cast<const float4*>(b + i), acc);
  return warp_sum(acc);
}
__device__ __forceinline__ float exact_sim(float dot, float nq, float ne) {
  return __fdiv_rn(dot, __fmul_rn(nq, ne));          // cache.py:291 / qre:138
}
__device__ __forceinline__ bool cand_better(float s0, uint64_t l0, float s1, uint64_t l1) {
  return s0 > s1 || (s0 == s1 && l0 < l1);           // (sim desc, lru_seq asc): cache.py:292
}
__device__ __forceinline__ qre_candidate_t empty_cand() {
  qre_candidate_t c;
  c.sim = -INFINITY;
  c.shard = 0;
  c.slot = -1;
  c.lru_seq = ~0ull;
  c.pred_conf = 0.0;
  return c;
}

// ---------------------------------------------------------------------------------------
// Row maintenance.  One warp per row.
// ---------------------------------------------------------------------------------------
struct WriteRowsArgs {
  long long n;
  const int64_t* slot;
  const float* rows;       // [n, rows_ld] or null
  long long rows_ld;
  const uint64_t* ctx;
  const int64_t* ts;
  const uint64_t* lru;
  const uint8_t* pdev;
  const double* pconf;
  int dim, dim_pad;
  float* rows_f32;
  __nv_bfloat16* rows_bf16;
  float* row_norm;
  RowMeta* meta;
  uint64_t* lru_out;
  uint8_t* pdev_out;
  double* pconf_out;
};

__global__ void write_rows_kernel(const WriteRowsArgs a) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= a.n) return;
  const long long slot = a.slot[w];
  if (a.rows != nullptr) {
    const float* src = a.rows + w * a.rows_ld;
    float* dst = a.rows_f32 + slot * a.dim_pad;
    float acc = 0.f;
    for (int i = lane * 4; i < a.dim_pad; i += 128) {
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < a.dim) x = *reinterpret_cast<const float4*>(src + i);
      *reinterpret_cast<float4*>(dst + i) = x;
      acc = lane_fma4(x, x, acc);
    }
    const float norm = __fsqrt_rn(warp_sum(acc));     // np.linalg.norm on fp32: sqrt32(sdot32)
    __nv_bfloat16* bdst = a.rows_bf16 + slot * a.dim_pad;
    const bool ok = norm >= kNormEps;
    for (int i = lane * 4; i < a.dim_pad; i += 128) {
      const float4 x = *reinterpret_cast<const float4*>(dst + i);
      __nv_bfloat162 lo, hi;
      lo.x = __float2bfloat16_rn(ok ? __fdiv_rn(x.x, norm) : 0.f);
      lo.y = __float2bfloat16_rn(ok ? __fdiv_rn(x.y, norm) : 0.f);
      hi.x = __float2bfloat16_rn(ok ? __fdiv_rn(x.z, norm) : 0.f);
      hi.y = __float2bfloat16_rn(ok ? __fdiv_rn(x.w, norm) : 0.f);
      *reinterpret_cast<__nv_bfloat162*>(bdst + i) = lo;
      *reinterpret_cast<__nv_bfloat162*>(bdst + i + 2) = hi;
    }
    if (lane == 0) a.row_norm[slot] = norm;
  }
  if (lane == 0) {
    if (a.ctx != nullptr) a.meta[slot].ctx = a.ctx[w];
    if (a.ts != nullptr) a.meta[slot].ts_us = a.ts[w];
    if (a.lru != nullptr) a.lru_out[slot] = a.lru[w];
    if (a.pdev != nullptr) a.pdev_out[slot] = a.pdev[w];
    if (a.pconf != nullptr) a.pconf_out[slot] = a.pconf[w];
  }
}

__global__ void tombstone_kernel(long long n, const int64_t* slot, RowMeta* meta, __nv_bfloat16* rows_bf16,
                                 float* row_norm, int dim_pad) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const long long s = slot[w];
  // zero the bf16 row so a deleted entry can never look like a candidate to the tensor-core scan
  uint2* b = reinterpret_cast<uint2*>(rows_bf16 + s * dim_pad);
  for (int i = lane; i < dim_pad / 4; i += 32) b[i] = make_uint2(0u, 0u);
  if (lane == 0) {
    meta[s].ctx = QRE_CTX_TOMBSTONE;
    row_norm[s] = 0.f;
  }
}

__global__ void gather_rows_kernel(long long n, const int64_t* slot, const float* rows_f32, int dim, int dim_pad,
                                   float* out) {
  const long long w = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n) return;
  const float* src = rows_f32 + slot[w] * dim_pad;
  for (int i = lane * 4; i < dim; i += 128)
    *reinterpret_cast<float4*>(out + w * dim + i) = *reinterpret_cast<const float4*>(src + i);
}

// Queries: fp32 copy (zero padded), |q|, and the L2-normalised bf16 operand of the scan.
__global__ void prep_queries_kernel(int m, int rows_total, const float* q, int dim, int dim_pad, float* q_f32,
                                    __nv_bfloat16* q_bf16, float* q_norm, const uint64_t* q_ctx_src,
                                    uint64_t* q_ctx_dst, uint32_t* counters, const uint32_t* wait_flags,
                                    int n_wait, uint32_t wait_epoch) {
  pdl_launch_dependents();      // head of the small-batch chain: let the scan's CTAs set up meanwhile
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (blockIdx.x == 0 && threadIdx.x < 4) counters[threadIdx.x] = 0u;   // spill bump, overflow counters
  if (wait_flags != nullptr) {
    // The batch's queries were uploaded in slices, one per rank, and stored into THIS GPU's input
    // buffer by the ranks' scatter kernels: wait for every rank's flag (bounded, like the merge).
    if (threadIdx.x < n_wait) {
      const volatile uint32_t* f = wait_flags + threadIdx.x * 32;
      const long long t0 = clock64();
      while (static_cast<int32_t>(*f - wait_epoch) < 0) {
        if (clock64() - t0 > 8000000000LL) {
          printf("qre: rank %d never delivered its query slice for epoch %u (saw %u)\n", threadIdx.x, wait_epoch, *f);
          __trap();
        }
      }
      __threadfence_system();
    }
    __syncthreads();
  }
  if (w >= rows_total) return;
  // (inputs may have been stored by peer GPUs: read them through L2, never from a stale L1 line)
  if (lane == 0 && w < m) q_ctx_dst[w] = __ldcg(q_ctx_src + w);  // private copy for the scan
  float* dst = q_f32 + static_cast<long long>(w) * dim_pad;
  __nv_bfloat16* bdst = q_bf16 + static_cast<long long>(w) * dim_pad;
  float acc = 0.f;
  for (int i = lane * 4; i < dim_pad; i += 128) {
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (w < m && i < dim) x = __ldcg(reinterpret_cast<const float4*>(q + static_cast<long long>(w) * dim + i));
    *reinterpret_cast<float4*>(dst + i) = x;
    acc = lane_fma4(x, x, acc);
  }
  const float norm = __fsqrt_rn(warp_sum(acc));
  const bool ok = norm >= kNormEps;
  for (int i = lane * 4; i < dim_pad; i += 128) {
    const float4 x = *reinterpret_cast<const float4*>(dst + i);
    __nv_bfloat162 lo, hi;
    lo.x = __float2bfloat16_rn(ok ? __fdiv_rn(x.x, norm) : 0.f);
    lo.y = __float2bfloat16_rn(ok ? __fdiv_rn(x.y, norm) : 0.f);
    hi.x = __float2bfloat16_rn(ok ? __fdiv_rn(x.z, norm) : 0.f);
    hi.y = __float2bfloat16_rn(ok ? __fdiv_rn(x.w, norm) : 0.f);
    *reinterpret_cast<__nv_bfloat162*>(bdst + i) = lo;
    *reinterpret_cast<__nv_bfloat162*>(bdst + i + 2) = hi;
  }
  if (lane == 0) q_norm[w] = norm;
}

__global__ void prep_centroids_kernel(int k, const float* c, int dim_pad, float* c_norm) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= k) return;
  const float* p = c + static_cast<long long>(w) * dim_pad;
  const float d = warp_dot(p, p, dim_pad, lane);
  if (lane == 0) c_norm[w] = __fsqrt_rn(d);
}

// ---------------------------------------------------------------------------------------
// Exact re-score of the tensor-core candidates.  One warp per query.
// ---------------------------------------------------------------------------------------
struct RescoreArgs {
  int m, splits, dim_pad, shard;
  float floor_v;
  double thr;
  const PartialEntry* partial;
  const uint32_t* spill_head;
  const SpillChunk* spill_pool;
  const float* q_f32;
  const float* q_norm;
  const uint64_t* q_ctx;
  const RowMeta* meta;
  long long now_us, ttl_us;
  const float* rows_f32;
  const float* row_norm;
  const uint64_t* lru;
  const uint8_t* pdev;
  const double* pconf;
  qre_candidate_t* cand;     // [m, kTopK]
  uint8_t* q_flags;          // [m]
  int32_t* ovf_count;
  int32_t* ovf_list;
};

struct CandList {
  float sim[kTopK];
  uint32_t row[kTopK];
  uint64_t lru[kTopK];
  int n;
};

__device__ __forceinline__ void list_insert(CandList& L, float sim, uint32_t row, uint64_t lru) {
  int i = L.n < kTopK ? L.n : kTopK - 1;
  if (L.n == kTopK && !cand_better(sim, lru, L.sim[i], L.lru[i])) return;
  while (i > 0 && cand_better(sim, lru, L.sim[i - 1], L.lru[i - 1])) {
    L.sim[i] = L.sim[i - 1];
    L.row[i] = L.row[i - 1];
    L.lru[i] = L.lru[i - 1];
    --i;
  }
  L.sim[i] = sim;
  L.row[i] = row;
  L.lru[i] = lru;
  if (L.n < kTopK) ++L.n;
}

// Exact score of one candidate row for the warp's query; lane 0 keeps the (sim desc, lru asc) list.
__device__ __forceinline__ void rescore_one(const RescoreArgs& a, CandList& L, const float* qv, float nq,
                                            uint32_t row, int lane) {
  const float ne = a.row_norm[row];
  const float dot = warp_dot(qv, a.rows_f32 + static_cast<long long>(row) * a.dim_pad, a.dim_pad, lane);
  if (ne >= kNormEps) {                                // cache.py:289-290
    const float sim = exact_sim(dot, nq, ne);
    if (static_cast<double>(sim) >= a.thr && sim > 0.f) {          // cache.py:292, best_sim starts at 0.0
      if (lane == 0) list_insert(L, sim, row, a.lru[row]);
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) rescore_kernel(const RescoreArgs a) {
  __shared__ CandList lists[kWarpsPerBlock];
  pdl_launch_dependents();
  pdl_wait();                   // no-ops unless launched as a link of the small-batch chain
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = blockIdx.x * kWarpsPerBlock + warp;
  if (q >= a.m) return;
  CandList& L = lists[warp];
  if (lane == 0) L.n = 0;
  __syncwarp();
  uint8_t flags = 0;
  const float nq = a.q_norm[q];
  bool ovf = false;
  if (nq < kNormEps) {
    flags |= QRE_FLAG_ZERO_NORM;                     // cache.py:284-285: return None
  } else {
    const PartialEntry* pe = a.partial + static_cast<long long>(q) * a.splits * kTopK;
    const uint32_t* heads = a.spill_head + static_cast<long long>(q) * a.splits;
    const int cnt = a.splits * kTopK;
    const uint64_t myctx = a.q_ctx[q];
    // The scan appended every score above its gate WITHOUT looking at the row (qre_scan_epilogue.cuh):
    // the row mask (context, TTL: cache.py:274-276) is applied here, 32 candidates at a time.
    auto valid = [&](const PartialEntry& e) {
      if (e.row == kNoRow || !(e.v >= a.floor_v)) return false;
      const RowMeta mt = a.meta[e.row];
      return mt.ctx == myctx && (a.now_us - mt.ts_us) <= a.ttl_us;
    };
    // Both passes walk the same entries: the inline slots of every range, then the chains.  Loads are
    // issued eight deep per lane before anything is looked at: with one query block the scan has
    // one row range per SM (148 ranges = 1,184 slots per query) and a walk that waits for each load
    // in turn costs ~15 us of a 200 us small-batch call (batch 1, p50 211 -> 194 us).
    auto walk = [&](auto&& visit) {
      for (int base = 0; base < cnt; base += 256) {
        PartialEntry e[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int i = base + j * 32 + lane;
          e[j].v = 0.f;
          e[j].row = kNoRow;
          if (i < cnt) e[j] = pe[i];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (base + j * 32 < cnt) visit(e[j]);
      }
      for (int base = 0; base < a.splits; base += 256) {
        uint32_t hc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int sidx = base + j * 32 + lane;
          hc[j] = sidx < a.splits ? heads[sidx] : kNoChunk;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          unsigned chains = __ballot_sync(kFull, hc[j] != kNoChunk && hc[j] != kChunkLost);
          while (chains) {
            const int src = __ffs(chains) - 1;
            chains &= chains - 1;
            uint32_t c = __shfl_sync(kFull, hc[j], src);
            while (c != kNoChunk) {
              const SpillChunk* ch = a.spill_pool + c;
              const uint32_t n = ch->n;
              PartialEntry e;
              e.v = 0.f;
              e.row = kNoRow;
              if (lane < static_cast<int>(n)) e = ch->e[lane];
              visit(e);
              c = ch->next;
            }
          }
        }
      }
    };
    for (int sidx = lane; sidx < a.splits; sidx += 32)
      if (heads[sidx] == kChunkLost) ovf = true;         // the spill pool ran out inside this range
    ovf = __any_sync(kFull, ovf);
    if (!ovf) {
      // pass 1: a* = best approximate score of any valid row that was appended
      float amax = -INFINITY;
      walk([&](const PartialEntry& e) {
        if (valid(e)) amax = fmaxf(amax, e.v);
      });
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(kFull, amax, o));
      if (amax != -INFINITY) {
        // pass 2: the exact winner has approx >= a* - 2 eps; nothing under the floor can hit
        const float cut = fmaxf(amax - 2.f * kApproxEps, a.floor_v);
        const float* qv = a.q_f32 + static_cast<long long>(q) * a.dim_pad;
        walk([&](const PartialEntry& e) {
          unsigned mask = __ballot_sync(kFull, e.v >= cut && valid(e));
          while (mask) {
            const int src = __ffs(mask) - 1;
            mask &= mask - 1;
            rescore_one(a, L, qv, nq, __shfl_sync(kFull, e.row, src), lane);
          }
        });
      }
    }
  }
  __syncwarp();
  if (lane < kTopK) {
    qre_candidate_t c = empty_cand();
    if (lane < L.n) {
      const uint32_t row = L.row[lane];
      c.sim = L.sim[lane];
      c.slot = row;
      c.lru_seq = L.lru[lane];
      c.shard = static_cast<uint32_t>(a.shard) | (static_cast<uint32_t>(a.pdev[row]) << 24);
      c.pred_conf = a.pconf[row];
    }
    a.cand[static_cast<long long>(q) * kTopK + lane] = c;
  }
  if (lane == 0) {
    if (ovf) {
      flags |= QRE_FLAG_OVERFLOW;
      const int idx = atomicAdd(a.ovf_count, 1);
      a.ovf_list[idx] = q;
    }
    a.q_flags[q] = flags;
  }
}

// ---------------------------------------------------------------------------------------
// Brute-force exact scan (fallback for overflowed queries, and the QRE_SCAN_EXACT mode).
// Up to kExactGroup queries per pass; each CTA owns a contiguous slice of rows, one warp per
// row, fp32 master rows streamed once (HBM-bound).
// ---------------------------------------------------------------------------------------
struct ExactArgs {
  int nq;                       // queries in this group (<= kExactGroup)
  int q_first;                  // index into the query list of the group's first query
  const int32_t* qlist;         // query indices (null = identity)
  long long n_rows;
  int dim_pad;
  double thr;
  long long now_us, ttl_us;
  const float* q_f32;
  const float* q_norm;
  const uint64_t* q_ctx;
  const float* rows_f32;
  const float* row_norm;
  const RowMeta* meta;
  const uint64_t* lru;
  ExactBest* partial;           // [kExactGroup, gridDim.x]
};

constexpr int kExactWarps = 8;
constexpr int kExactMaxIt = 8;  // dim_pad <= 1024

struct ExactSmem {
  float nq[kExactGroup];
  uint64_t ctx[kExactGroup];
  ExactBest best[kExactWarps][kExactGroup];
};

// One CTA's share of a brute-force pass: stage the group's queries, scan rows [r0, r1) of this
// CTA's slice, leave the per-query winners in a.partial[query][blockIdx.x].
__device__ __forceinline__ void exact_scan_body(const ExactArgs& a, float* qs, ExactSmem& sm) {
  float* s_nq = sm.nq;
  uint64_t* s_ctx = sm.ctx;
  ExactBest (*s_best)[kExactGroup] = sm.best;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = 0; j < a.nq; ++j) {
    const int q = a.qlist ? a.qlist[a.q_first + j] : a.q_first + j;
    for (int i = threadIdx.x; i < a.dim_pad; i += blockDim.x)
      qs[j * a.dim_pad + i] = a.q_f32[static_cast<long long>(q) * a.dim_pad + i];
    if (threadIdx.x == 0) {
      s_nq[j] = a.q_norm[q];
      s_ctx[j] = a.q_ctx[q];
    }
  }
  __syncthreads();
  const long long r0 = a.n_rows * blockIdx.x / gridDim.x, r1 = a.n_rows * (blockIdx.x + 1) / gridDim.x;
  float bsim = -INFINITY;
  uint32_t brow = kNoRow;
  uint64_t blru = ~0ull;
  const int nit = (a.dim_pad + 127) / 128;
  for (long long row = r0 + warp; row < r1; row += kExactWarps) {
    const RowMeta mt = a.meta[row];
    if (mt.ctx == QRE_CTX_TOMBSTONE || (a.now_us - mt.ts_us) > a.ttl_us) continue;   // cache.py:275-276
    const float ne = a.row_norm[row];
    if (ne < kNormEps) continue;                                                      // cache.py:289
    const bool mine = lane < a.nq && s_ctx[lane] == mt.ctx && s_nq[lane] >= kNormEps; // cache.py:274,284
    if (!__any_sync(kFull, mine)) continue;
    float4 e[kExactMaxIt];
    const float* rp = a.rows_f32 + row * a.dim_pad;
#pragma unroll
    for (int it = 0; it < kExactMaxIt; ++it) {
      const int i = lane * 4 + it * 128;
      e[it] = (it < nit && i < a.dim_pad) ? *reinterpret_cast<const float4*>(rp + i)
                                          : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float mydot = 0.f;
    for (int j = 0; j < a.nq; ++j) {
      float acc = 0.f;
#pragma unroll
      for (int it = 0; it < kExactMaxIt; ++it) {
        const int i = lane * 4 + it * 128;
        if (it < nit && i < a.dim_pad)
          acc = lane_fma4(*reinterpret_cast<const float4*>(qs + j * a.dim_pad + i), e[it], acc);
      }
      acc = warp_sum(acc);
      if (lane == j) mydot = acc;
    }
    if (mine) {
      const float sim = exact_sim(mydot, s_nq[lane], ne);
      if (static_cast<double>(sim) >= a.thr && sim > 0.f) {
        const uint64_t l = a.lru[row];
        if (cand_better(sim, l, bsim, blru)) {
          bsim = sim;
          blru = l;
          brow = static_cast<uint32_t>(row);
        }
      }
    }
  }
  if (lane < kExactGroup) s_best[warp][lane] = ExactBest{bsim, brow, blru};
  __syncthreads();
  if (warp == 0 && lane < a.nq) {
    ExactBest b = s_best[0][lane];
    for (int w = 1; w < kExactWarps; ++w) {
      const ExactBest c = s_best[w][lane];
      if (c.row != kNoRow && (b.row == kNoRow || cand_better(c.sim, c.lru, b.sim, b.lru))) b = c;
    }
    a.partial[static_cast<long long>(lane) * gridDim.x + blockIdx.x] = b;
  }
}

__global__ void __launch_bounds__(kExactWarps * 32) exact_scan_kernel(const ExactArgs a) {
  extern __shared__ float qs[];                     // [kExactGroup, dim_pad]
  __shared__ ExactSmem sm;
  exact_scan_body(a, qs, sm);
}

// Cross-CTA reduction of one group's partial winners -> cand[q][0] (warp j handles query j).
__device__ __forceinline__ void exact_reduce_body(int nq, int q_first, const int32_t* qlist, int grid,
                                                  const ExactBest* partial, int shard, const uint8_t* pdev,
                                                  const double* pconf, const float* q_norm,
                                                  qre_candidate_t* cand, uint8_t* q_flags, uint8_t extra_flag) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp >= nq) return;
  const int q = qlist ? qlist[q_first + warp] : q_first + warp;
  ExactBest b{-INFINITY, kNoRow, ~0ull};
  for (int i = lane; i < grid; i += 32) {
    const ExactBest* pp = &partial[static_cast<long long>(warp) * grid + i];
    ExactBest c;                                    // may have been written by another CTA of this launch
    c.sim = __ldcg(&pp->sim);
    c.row = __ldcg(&pp->row);
    c.lru = __ldcg(&pp->lru);
    if (c.row != kNoRow && (b.row == kNoRow || cand_better(c.sim, c.lru, b.sim, b.lru))) b = c;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ExactBest c;
    c.sim = __shfl_xor_sync(kFull, b.sim, o);
    c.row = __shfl_xor_sync(kFull, b.row, o);
    c.lru = __shfl_xor_sync(kFull, b.lru, o);
    if (c.row != kNoRow && (b.row == kNoRow || cand_better(c.sim, c.lru, b.sim, b.lru))) b = c;
  }
  if (lane < kTopK) {
    qre_candidate_t c = empty_cand();
    if (lane == 0 && b.row != kNoRow) {
      c.sim = b.sim;
      c.slot = b.row;
      c.lru_seq = b.lru;
      c.shard = static_cast<uint32_t>(shard) | (static_cast<uint32_t>(pdev[b.row]) << 24);
      c.pred_conf = pconf[b.row];
    }
    cand[static_cast<long long>(q) * kTopK + lane] = c;
  }
  if (lane == 0) q_flags[q] = extra_flag | (q_norm[q] < kNormEps ? QRE_FLAG_ZERO_NORM : 0);
}

__global__ void exact_reduce_kernel(int nq, int q_first, const int32_t* qlist, int grid, const ExactBest* partial,
                                    int shard, const uint8_t* pdev, const double* pconf, const float* q_norm,
                                    qre_candidate_t* cand, uint8_t* q_flags, uint8_t extra_flag) {
  exact_reduce_body(nq, q_first, qlist, grid, partial, shard, pdev, pconf, q_norm, cand, q_flags, extra_flag);
}

// Device-driven overflow settlement: ONE launch after rescore_kernel, no host round trip.  Reads
// the overflow counter the re-score pass left on the device; exits at once when it is zero (the
// normal case: candidates that do not fit a list are chained, so a query overflows only when the
// spill pool itself ran out); otherwise runs the brute-force pass over groups of 8 overflowed
// queries, the last CTA of each group (ticket counter) doing the reduction.  The ticket / partial
// arrays hold max_batch / 8 groups, so every query of a batch can be settled here and
// *unresolved is always left at 0.
struct SettleArgs {
  ExactArgs base;                 // nq / q_first / partial filled per group
  const int32_t* ovf_count;
  const int32_t* ovf_list;
  int32_t* unresolved;
  unsigned* tickets;              // [max_groups], zero between launches
  ExactBest* partial;             // [max_groups][kExactGroup][gridDim.x]
  int max_groups;                 // ceil(max_batch / kExactGroup): enough for every query of a batch
  int shard;
  const uint8_t* pdev;
  const double* pconf;
  qre_candidate_t* cand;
  uint8_t* q_flags;
};

__global__ void __launch_bounds__(kExactWarps * 32) overflow_settle_kernel(const SettleArgs sa) {
  extern __shared__ float qs[];
  __shared__ ExactSmem sm;
  __shared__ unsigned s_ticket;
  pdl_launch_dependents();
  pdl_wait();
  const int n_ovf = *sa.ovf_count;
  if (n_ovf <= 0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) *sa.unresolved = 0;
    return;
  }
  const int groups_all = (n_ovf + kExactGroup - 1) / kExactGroup;
  const int groups = groups_all < sa.max_groups ? groups_all : sa.max_groups;
  for (int g = 0; g < groups; ++g) {
    ExactArgs a = sa.base;
    a.q_first = g * kExactGroup;
    a.nq = n_ovf - a.q_first < kExactGroup ? n_ovf - a.q_first : kExactGroup;
    a.qlist = sa.ovf_list;
    a.partial = sa.partial + static_cast<long long>(g) * kExactGroup * gridDim.x;
    __syncthreads();
    exact_scan_body(a, qs, sm);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_ticket = atomicAdd(&sa.tickets[g], 1u);
    __syncthreads();
    if (s_ticket == gridDim.x - 1) {          // last CTA of this group: all partials are visible
      __threadfence();
      exact_reduce_body(a.nq, a.q_first, a.qlist, gridDim.x, a.partial, sa.shard, sa.pdev, sa.pconf, a.q_norm,
                        sa.cand, sa.q_flags, QRE_FLAG_OVERFLOW);
      if (threadIdx.x == 0) {
        sa.tickets[g] = 0u;
        if (g == groups - 1) *sa.unresolved = n_ovf - groups * kExactGroup > 0 ? n_ovf - groups * kExactGroup : 0;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Merge of per-shard candidates + the routing epilogue.  One warp per query.
// ---------------------------------------------------------------------------------------
struct MergeArgs {
  int m, n_shards, dim_pad, n_centroids, do_route;
  const qre_candidate_t* cand_all;   // [n_shards, m, kTopK]
  const float* q_f32;
  const float* q_norm;
  const uint8_t* q_flags;
  const float* centroids;
  const float* cent_norm;
  const uint8_t* cent_owner;
  const double* perf_health;
  const int32_t* tok;
  const uint8_t* heur_dev;
  const double* heur_conf;
  qre_params_t p;
  qre_result_t out;
  int q0, q1;                        // queries [q0, q1) are handled by this launch (sliced results: one rank's share)
  long long shard_stride;            // records between consecutive shards' lists in cand_all
  const uint32_t* wait_flags;        // peer exchange: flags[shard * 32] must equal wait_epoch first
  uint32_t wait_epoch;
};

constexpr int kMergeMax = 128;     // n_shards * kTopK <= 128 (up to 16 shards)
constexpr int kMergePerLane = kMergeMax / 32;

// The routing epilogue of one query (one warp): SemanticRouter cosine to every centroid with
// Sim[dev] = max over owned centroids, TokenBasedRouter, HybridRouter.  Depends on the query and
// its host features only — not on the cache scan — so the pipeline runs it at preparation time.
__device__ __forceinline__ void route_query(const MergeArgs& a, int q, int lane, float* ds, double* S) {
  const int D = a.p.n_devices;
  for (int d = lane; d < D; d += 32) ds[d] = -INFINITY;
  __syncwarp();
  const float nq = a.q_norm[q];
  const float* qv = a.q_f32 + static_cast<long long>(q) * a.dim_pad;
  // The query row stays in registers (dim_pad <= 1024 -> at most 8 float4 per lane) and the
  // centroid loop is unrolled by 4 so the loads of independent centroids overlap; the summation
  // order per centroid is warp_dot's (every exact score in the library uses it).
  float4 qr[8];
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int i = lane * 4 + it * 128;
    qr[it] = i < a.dim_pad ? *reinterpret_cast<const float4*>(qv + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int k0 = 0; k0 < a.n_centroids; k0 += 4) {
    float dot[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      float acc = 0.f;
      if (k0 + u < a.n_centroids) {
        const float* cv = a.centroids + static_cast<long long>(k0 + u) * a.dim_pad;
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int i = lane * 4 + it * 128;
          if (i < a.dim_pad) acc = lane_fma4(qr[it], __ldg(reinterpret_cast<const float4*>(cv + i)), acc);
        }
      }
      dot[u] = acc;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) dot[u] = warp_sum(dot[u]);
    if (lane == 0) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + u;
        if (k >= a.n_centroids) break;
        const float nc = a.cent_norm[k];
        const float sim = (nq < kNormEps || nc < kNormEps) ? 0.f : exact_sim(dot[u], nq, nc);   // qre:135-138
        if (a.out.centroid_sim) a.out.centroid_sim[static_cast<long long>(q) * a.n_centroids + k] = sim;
        const int o = a.cent_owner[k];
        if (sim > ds[o]) ds[o] = sim;
      }
    }
  }
  if (lane != 0) return;

  // TokenBasedRouter (qre:98-99)
  const int tok = a.tok[q], thr = a.p.token_threshold;
  const int tok_dev = (tok > thr && D > 1) ? 1 : 0;   // "orin" = device 1; a single-device set-up has only device 0
  const int den = thr > 1 ? thr : 1;
  const int diff = tok > thr ? tok - thr : thr - tok;
  double tok_conf = static_cast<double>(diff) / static_cast<double>(den);
  if (tok_conf > 1.0) tok_conf = 1.0;

  // SemanticRouter.route (qre:185-213), generalised to D devices
  int sem_dev, sem_method;
  double sem_conf;
  {
    bool all_low = true;
    int top = -1;
    for (int d = 0; d < D; ++d) {
      if (ds[d] == -INFINITY) continue;
      if (!(static_cast<double>(ds[d]) < a.p.sem_min_similarity)) all_low = false;
      if (top < 0 || ds[d] > ds[top]) top = d;
    }
    double margin = INFINITY;
    if (top >= 0) {
      float second = -INFINITY;
      bool any = false;
      for (int d = 0; d < D; ++d)
        if (d != top && ds[d] != -INFINITY) {
          any = true;
          if (ds[d] > second) second = ds[d];
        }
      if (any) margin = fabs(__dsub_rn(static_cast<double>(ds[top]), static_cast<double>(second)));
    }
    if (all_low) {
      sem_dev = tok_dev;
      sem_conf = __dmul_rn(tok_conf, 0.5);
      sem_method = QRE_SEM_IRRELEVANT;
    } else if (margin < a.p.sem_margin_threshold) {
      sem_dev = tok_dev;
      sem_conf = margin;
      sem_method = QRE_SEM_AMBIGUOUS;
    } else {
      sem_dev = top;
      const double c = margin / 0.2;
      sem_conf = c < 1.0 ? c : 1.0;
      sem_method = QRE_SEM_SEMANTIC;
    }
  }

  // HybridRouter.route (qre:389-407): S[dev] += w * conf in the order token, semantic, heuristic
  for (int d = 0; d < D; ++d) S[d] = 0.0;
  S[tok_dev] = __dadd_rn(S[tok_dev], __dmul_rn(a.p.w_token, tok_conf));
  S[sem_dev] = __dadd_rn(S[sem_dev], __dmul_rn(a.p.w_semantic, sem_conf));
  int hd = a.heur_dev[q];
  if (hd >= D) hd = D - 1;                            // header contract: heur_dev < n_devices; never index past S[D)
  S[hd] = __dadd_rn(S[hd], __dmul_rn(a.p.w_heuristic, a.heur_conf[q]));
  if (a.p.w_perf != 0.0)
    for (int d = 0; d < D; ++d) S[d] = __dadd_rn(S[d], __dmul_rn(a.p.w_perf, a.perf_health[d]));
  int top = 0;
  for (int d = 1; d < D; ++d)
    if (S[d] > S[top]) top = d;                                    // ties -> lowest index (qre:401)
  double second = -INFINITY, total = 0.0;
  for (int d = 0; d < D; ++d) {
    if (d != top && S[d] > second) second = S[d];
    total = __dadd_rn(total, S[d]);
  }
  const double margin = D > 1 ? __dsub_rn(S[top], second) : S[top];
  double conf = total > 1e-12 ? margin / total : 0.5;              // qre:407
  conf = conf < 0.0 ? 0.0 : (conf > 1.0 ? 1.0 : conf);

  if (a.out.tok_dev) a.out.tok_dev[q] = static_cast<uint8_t>(tok_dev);
  if (a.out.tok_conf) a.out.tok_conf[q] = tok_conf;
  if (a.out.sem_dev) a.out.sem_dev[q] = static_cast<uint8_t>(sem_dev);
  if (a.out.sem_conf) a.out.sem_conf[q] = sem_conf;
  if (a.out.sem_method) a.out.sem_method[q] = static_cast<uint8_t>(sem_method);
  if (a.out.dev) a.out.dev[q] = static_cast<uint8_t>(top);
  if (a.out.conf) a.out.conf[q] = conf;
  if (a.out.score)
    for (int d = 0; d < D; ++d) a.out.score[static_cast<long long>(q) * D + d] = S[d];
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32) route_only_kernel(const MergeArgs a) {
  __shared__ float s_dsim[kWarpsPerBlock][QRE_MAX_DEVICES];
  __shared__ double s_S[kWarpsPerBlock][QRE_MAX_DEVICES];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = a.q0 + blockIdx.x * kWarpsPerBlock + warp;
  if (q >= a.q1) return;
  route_query(a, q, lane, s_dsim[warp], s_S[warp]);
}

// Candidates live in registers (lane l holds candidates l, l+32, l+64, l+96): no shared-memory
// staging, so blocks of this kernel fit beside a resident scan CTA.
__global__ void __launch_bounds__(kWarpsPerBlock * 32) merge_route_kernel(const MergeArgs a) {
  __shared__ float s_dsim[kWarpsPerBlock][QRE_MAX_DEVICES];
  __shared__ double s_S[kWarpsPerBlock][QRE_MAX_DEVICES];
  pdl_launch_dependents();
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = a.q0 + blockIdx.x * kWarpsPerBlock + warp;
  if (a.wait_flags != nullptr) {
    // Peer exchange: every rank's publish kernel stores its candidates into THIS GPU's buffer over
    // NVLink and then raises its flag to the batch's epoch.  The publishers run on other GPUs, so
    // this wait never depends on work queued behind it on this GPU.  Bounded: a dead peer must
    // abort the job, not hang the box.
    if (threadIdx.x < a.n_shards) {
      const volatile uint32_t* f = a.wait_flags + threadIdx.x * 32;
      const long long t0 = clock64();
      // signed distance: a flag that has already moved past this epoch cannot hang the wait
      while (static_cast<int32_t>(*f - a.wait_epoch) < 0) {
        if (clock64() - t0 > 8000000000LL) {
          printf("qre: peer %d never published epoch %u (saw %u)\n", threadIdx.x, a.wait_epoch, *f);
          __trap();
        }
      }
      __threadfence_system();
    }
    __syncthreads();
  }
  if (q >= a.q1) return;

  // ---- merge: n_shards * kTopK candidates, comparator (sim desc, lru asc) ----
  const int n_cand = a.n_shards * kTopK;
  qre_candidate_t c[kMergePerLane];
#pragma unroll
  for (int j = 0; j < kMergePerLane; ++j) {
    const int i = lane + 32 * j;
    c[j] = empty_cand();
    if (i < n_cand) {
      const int r = i / kTopK, k = i - r * kTopK;
      const int4* src = reinterpret_cast<const int4*>(a.cand_all + static_cast<long long>(r) * a.shard_stride +
                                                      static_cast<long long>(q) * kTopK + k);
      int4* dst = reinterpret_cast<int4*>(&c[j]);    // read through L2: peers may have written it
      dst[0] = __ldcg(src);
      dst[1] = __ldcg(src + 1);
    }
  }
  // Most queries have zero or one live candidate, and callers that do not ask for the merged
  // list only need the winner: run only as many selection rounds as can produce something.
  int n_valid = 0;
#pragma unroll
  for (int j = 0; j < kMergePerLane; ++j) n_valid += c[j].slot >= 0 ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n_valid += __shfl_xor_sync(kFull, n_valid, o);
  const int want = a.out.cand != nullptr ? kTopK : 1;
  const int rounds = n_valid < want ? n_valid : want;
  if (a.out.cand != nullptr && lane >= rounds && lane < kTopK)
    a.out.cand[static_cast<long long>(q) * kTopK + lane] = empty_cand();
  qre_candidate_t best0 = empty_cand();
  for (int k = 0; k < rounds; ++k) {
    int bi = -1;
    float bs = -INFINITY;
    uint64_t bl = ~0ull;
#pragma unroll
    for (int j = 0; j < kMergePerLane; ++j) {
      if (c[j].slot >= 0 && (bi < 0 || cand_better(c[j].sim, c[j].lru_seq, bs, bl))) {
        bi = lane + 32 * j; bs = c[j].sim; bl = c[j].lru_seq;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float os = __shfl_xor_sync(kFull, bs, o);
      const uint64_t ol = __shfl_xor_sync(kFull, bl, o);
      const int oi = __shfl_xor_sync(kFull, bi, o);
      if (oi >= 0 && (bi < 0 || cand_better(os, ol, bs, bl) || (os == bs && ol == bl && oi < bi))) {
        bi = oi; bs = os; bl = ol;
      }
    }
    // the winner's owner hands the record to lane 0 and retires it
    const int owner = bi & 31, wj = bi >> 5;
    qre_candidate_t win = empty_cand();
#pragma unroll
    for (int j = 0; j < kMergePerLane; ++j) {
      if (j == wj) {
        win.sim = __shfl_sync(kFull, c[j].sim, owner);
        win.shard = __shfl_sync(kFull, c[j].shard, owner);
        win.slot = __shfl_sync(kFull, c[j].slot, owner);
        win.lru_seq = __shfl_sync(kFull, c[j].lru_seq, owner);
        win.pred_conf = __shfl_sync(kFull, c[j].pred_conf, owner);
        if (lane == owner) c[j].slot = -1;
      }
    }
    if (k == 0) best0 = win;
    if (lane == 0 && a.out.cand != nullptr) a.out.cand[static_cast<long long>(q) * kTopK + k] = win;
  }
  if (lane == 0) {
    const bool hit = best0.slot >= 0;
    if (a.out.hit_slot) a.out.hit_slot[q] = hit ? best0.slot : -1;
    if (a.out.hit_sim) a.out.hit_sim[q] = hit ? best0.sim : 0.f;
    if (a.out.hit_pred_dev) a.out.hit_pred_dev[q] = hit ? static_cast<uint8_t>(best0.shard >> 24) : 0;
    if (a.out.hit_pred_conf) a.out.hit_pred_conf[q] = hit ? best0.pred_conf : 0.0;
    if (a.out.hit_shard) a.out.hit_shard[q] = hit ? (best0.shard & 0xFFFFFFu) : 0u;
    if (a.out.flags) a.out.flags[q] = a.q_flags[q];
  }
  if (!a.do_route) return;
  route_query(a, q, lane, s_dsim[warp], s_S[warp]);
}

// ---------------------------------------------------------------------------------------
// launch wrappers
// ---------------------------------------------------------------------------------------
static inline unsigned blocks_for_warps(long long warps, int threads) {
  const long long per = threads / 32;
  return static_cast<unsigned>((warps + per - 1) / per);
}

static int write_rows(DeviceState& s, long long n, const float* rows, long long ld, const int64_t* slot,
                      const uint64_t* ctx, const int64_t* ts, const uint64_t* lru, const uint8_t* pdev,
                      const double* pconf, cudaStream_t st) {
  if (n <= 0) return 0;
  WriteRowsArgs a;
  a.n = n; a.slot = slot; a.rows = rows; a.rows_ld = ld; a.ctx = ctx; a.ts = ts; a.lru = lru;
  a.pdev = pdev; a.pconf = pconf; a.dim = s.dim; a.dim_pad = s.dim_pad; a.rows_f32 = s.rows_f32;
  a.rows_bf16 = s.rows_bf16; a.row_norm = s.row_norm; a.meta = s.meta; a.lru_out = s.lru;
  a.pdev_out = s.pred_dev; a.pconf_out = s.pred_conf;
  write_rows_kernel<<<blocks_for_warps(n, 256), 256, 0, st>>>(a);
  return 1;
}

int launch_append(DeviceState& s, long long n, const float* rows_dev, long long rows_ld, const int64_t* slot_dev,
                  const uint64_t* ctx, const int64_t* ts, const uint64_t* lru, const uint8_t* pdev,
                  const double* pconf, cudaStream_t st) {
  return write_rows(s, n, rows_dev, rows_ld, slot_dev, ctx, ts, lru, pdev, pconf, st);
}
int launch_update_meta(DeviceState& s, long long n, const int64_t* slot_dev, const uint64_t* ctx,
                       const int64_t* ts, const uint64_t* lru, const uint8_t* pdev, const double* pconf,
                       cudaStream_t st) {
  return write_rows(s, n, nullptr, 0, slot_dev, ctx, ts, lru, pdev, pconf, st);
}
int launch_tombstone(DeviceState& s, long long n, const int64_t* slot_dev, cudaStream_t st) {
  if (n <= 0) return 0;
  tombstone_kernel<<<blocks_for_warps(n, 256), 256, 0, st>>>(n, slot_dev, s.meta, s.rows_bf16, s.row_norm,
                                                            s.dim_pad);
  return 1;
}
int launch_gather_rows(DeviceState& s, long long n, const int64_t* slot_dev, float* out, cudaStream_t st) {
  if (n <= 0) return 0;
  gather_rows_kernel<<<blocks_for_warps(n, 256), 256, 0, st>>>(n, slot_dev, s.rows_f32, s.dim, s.dim_pad, out);
  return 1;
}
int launch_prep_queries(DeviceState& s, int m, const float* q_dev, const uint64_t* q_ctx_src, cudaStream_t st,
                        const uint32_t* wait_flags, int n_wait, uint32_t wait_epoch) {
  const int rows_total = ((m + kBM - 1) / kBM) * kBM;
  DeviceState::Batch& b = s.cur();
  prep_queries_kernel<<<blocks_for_warps(rows_total, 256), 256, 0, st>>>(
      m, rows_total, q_dev, s.dim, s.dim_pad, b.q_f32, b.q_bf16, b.q_norm, q_ctx_src, b.q_ctx, b.counters,
      wait_flags, n_wait, wait_epoch);
  return 1;
}

// Input exchange of the host-buffer path on a row-sharded cache: rank r uploaded queries
// [q0, q1) only; this kernel stores them (and their context ids) into every rank's input
// buffer over NVLink (plain stores) and then raises rank r's flag there.
struct ScatterArgs {
  int world, q0, q1, dim;
  const float* src_q;            // [m, dim] layout; rows [q0, q1) valid
  const uint64_t* src_ctx;       // [m]
  float* dst_q[16];
  uint64_t* dst_ctx[16];
  uint32_t* flag[16];
  uint32_t epoch;
  unsigned* ticket;
};

__global__ void __launch_bounds__(256) scatter_inputs_kernel(const ScatterArgs a) {
  __shared__ unsigned s_ticket;
  const long long n4 = static_cast<long long>(a.q1 - a.q0) * a.dim / 4;     // dim % 8 == 0
  const float4* src = reinterpret_cast<const float4*>(a.src_q + static_cast<long long>(a.q0) * a.dim);
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  for (int p = 0; p < a.world; ++p) {
    float4* d = reinterpret_cast<float4*>(a.dst_q[p] + static_cast<long long>(a.q0) * a.dim);
    for (long long i = tid; i < n4; i += nth) d[i] = src[i];
    for (long long i = a.q0 + tid; i < a.q1; i += nth) a.dst_ctx[p][i] = a.src_ctx[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(a.ticket, 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence_system();
  if (threadIdx.x < a.world) *reinterpret_cast<volatile uint32_t*>(a.flag[threadIdx.x]) = a.epoch;
  if (threadIdx.x == 0) *a.ticket = 0u;
}

int launch_scatter_inputs(int world, int q0, int q1, int dim, const float* src_q, const uint64_t* src_ctx,
                          float* const* dst_q, uint64_t* const* dst_ctx, uint32_t* const* flag, uint32_t epoch,
                          unsigned* ticket, cudaStream_t st) {
  ScatterArgs a;
  a.world = world; a.q0 = q0; a.q1 = q1; a.dim = dim; a.src_q = src_q; a.src_ctx = src_ctx; a.epoch = epoch;
  a.ticket = ticket;
  for (int p = 0; p < world; ++p) { a.dst_q[p] = dst_q[p]; a.dst_ctx[p] = dst_ctx[p]; a.flag[p] = flag[p]; }
  scatter_inputs_kernel<<<16, 256, 0, st>>>(a);
  return 1;
}

int launch_prep_centroids(DeviceState& s, int k, cudaStream_t st) {
  prep_centroids_kernel<<<blocks_for_warps(k, 128), 128, 0, st>>>(k, s.centroids, s.dim_pad, s.cent_norm);
  return 1;
}

int launch_rescore(DeviceState& s, int m, long long n_rows, const qre_params_t& p, int shard,
                   qre_candidate_t* cand, cudaStream_t st) {
  const ScanGeometry g = plan_for(s, m, n_rows, p);
  const DeviceState::Batch& b = s.cur();
  RescoreArgs a;
  a.m = m; a.splits = g.splits; a.dim_pad = s.dim_pad; a.shard = shard;
  const double thr = p.similarity_threshold > 0.0 ? p.similarity_threshold : 0.0;
  a.floor_v = static_cast<float>(thr) - kApproxEps;
  a.thr = p.similarity_threshold;
  a.partial = b.partial; a.spill_head = b.spill_head; a.spill_pool = b.spill_pool;
  a.q_f32 = b.q_f32; a.q_norm = b.q_norm; a.rows_f32 = s.rows_f32;
  a.q_ctx = b.q_ctx; a.meta = s.meta; a.now_us = p.now_us; a.ttl_us = p.ttl_us;
  a.row_norm = s.row_norm; a.lru = s.lru; a.pdev = s.pred_dev; a.pconf = s.pred_conf; a.cand = cand;
  a.q_flags = b.q_flags;
  a.ovf_count = b.ovf_count(); a.ovf_list = b.ovf_list;
  launch_chain(rescore_kernel, dim3((m + kWarpsPerBlock - 1) / kWarpsPerBlock), dim3(kWarpsPerBlock * 32), 0, st,
               s.pdl, a);
  return 1;
}

static ExactArgs exact_args(const DeviceState& s, long long n_rows, const qre_params_t& p) {
  const DeviceState::Batch& b = s.cur();
  ExactArgs a;
  a.nq = 0; a.q_first = 0; a.qlist = nullptr; a.n_rows = n_rows; a.dim_pad = s.dim_pad;
  a.thr = p.similarity_threshold; a.now_us = p.now_us; a.ttl_us = p.ttl_us; a.q_f32 = b.q_f32;
  a.q_norm = b.q_norm; a.q_ctx = b.q_ctx; a.rows_f32 = s.rows_f32; a.row_norm = s.row_norm; a.meta = s.meta;
  a.lru = s.lru; a.partial = nullptr;
  return a;
}

int launch_exact_scan(DeviceState& s, int m, long long n_rows, const int32_t* qlist_dev,
                      int nq_total, bool mark_overflow, const qre_params_t& p, int shard, qre_candidate_t* cand,
                      cudaStream_t st) {
  (void)m;
  const size_t smem = static_cast<size_t>(kExactGroup) * s.dim_pad * sizeof(float);   // <= 32 KB
  int launches = 0;
  const DeviceState::Batch& b = s.cur();
  for (int g0 = 0; g0 < nq_total; g0 += kExactGroup) {
    ExactArgs a = exact_args(s, n_rows, p);
    a.nq = nq_total - g0 < kExactGroup ? nq_total - g0 : kExactGroup;
    a.q_first = g0; a.qlist = qlist_dev; a.partial = s.exact_partial;
    exact_scan_kernel<<<s.exact_grid, kExactWarps * 32, smem, st>>>(a);
    exact_reduce_kernel<<<1, kExactGroup * 32, 0, st>>>(a.nq, g0, qlist_dev, s.exact_grid, s.exact_partial, shard,
                                                       s.pred_dev, s.pred_conf, b.q_norm, cand, b.q_flags,
                                                       mark_overflow ? QRE_FLAG_OVERFLOW : 0);
    launches += 2;
  }
  return launches;
}

int launch_overflow_settle(DeviceState& s, long long n_rows, const qre_params_t& p, int shard,
                           qre_candidate_t* cand, cudaStream_t st) {
  const DeviceState::Batch& b = s.cur();
  SettleArgs sa;
  sa.base = exact_args(s, n_rows, p);
  sa.base.qlist = b.ovf_list;
  sa.ovf_count = b.ovf_count(); sa.ovf_list = b.ovf_list; sa.unresolved = b.ovf_unresolved();
  sa.tickets = s.settle_tickets; sa.partial = s.settle_partial; sa.max_groups = s.settle_groups;
  sa.shard = shard; sa.pdev = s.pred_dev;
  sa.pconf = s.pred_conf; sa.cand = cand; sa.q_flags = b.q_flags;
  const size_t smem = static_cast<size_t>(kExactGroup) * s.dim_pad * sizeof(float);
  launch_chain(overflow_settle_kernel, dim3(s.exact_grid), dim3(kExactWarps * 32), smem, st, s.pdl, sa);
  return 1;
}

int launch_merge_route(DeviceState& s, int m, int n_shards, const qre_candidate_t* cand_all, const int32_t* tok,
                       const uint8_t* heur_dev, const double* heur_conf, const qre_params_t& p,
                       const qre_result_t& out, cudaStream_t st, long long shard_stride, const uint32_t* wait_flags,
                       uint32_t wait_epoch, int q0, int q1) {
  MergeArgs a;
  a.m = m; a.n_shards = n_shards; a.dim_pad = s.dim_pad; a.n_centroids = s.n_centroids;
  a.do_route = p.do_route; a.cand_all = cand_all; a.q_f32 = s.cur().q_f32; a.q_norm = s.cur().q_norm;
  a.q_flags = s.cur().q_flags;
  a.centroids = s.centroids; a.cent_norm = s.cent_norm; a.cent_owner = s.cent_owner;
  a.perf_health = s.perf_health; a.tok = tok; a.heur_dev = heur_dev; a.heur_conf = heur_conf; a.p = p;
  a.out = out;
  a.shard_stride = shard_stride > 0 ? shard_stride : static_cast<long long>(m) * kTopK;
  a.wait_flags = wait_flags;
  a.wait_epoch = wait_epoch;
  a.q0 = q0 < 0 ? 0 : q0;
  a.q1 = q1 < 0 || q1 > m ? m : q1;
  const int nq = a.q1 - a.q0 > 0 ? a.q1 - a.q0 : 1;   // an empty slice still launches (it may have to wait for flags)
  launch_chain(merge_route_kernel, dim3((nq + kWarpsPerBlock - 1) / kWarpsPerBlock), dim3(kWarpsPerBlock * 32), 0, st,
               s.pdl, a);
  return 1;
}

int launch_route_only(DeviceState& s, int m, const int32_t* tok, const uint8_t* heur_dev, const double* heur_conf,
                      const qre_params_t& p, const qre_result_t& out, cudaStream_t st, int q0, int q1) {
  MergeArgs a;
  a.m = m; a.n_shards = 0; a.dim_pad = s.dim_pad; a.n_centroids = s.n_centroids;
  a.do_route = 1; a.cand_all = nullptr; a.q_f32 = s.cur().q_f32; a.q_norm = s.cur().q_norm;
  a.q_flags = s.cur().q_flags;
  a.centroids = s.centroids; a.cent_norm = s.cent_norm; a.cent_owner = s.cent_owner;
  a.perf_health = s.perf_health; a.tok = tok; a.heur_dev = heur_dev; a.heur_conf = heur_conf; a.p = p;
  a.out = out;
  a.shard_stride = 0; a.wait_flags = nullptr; a.wait_epoch = 0;
  a.q0 = q0 < 0 ? 0 : q0;
  a.q1 = q1 < 0 || q1 > m ? m : q1;
  if (a.q1 <= a.q0) return 0;
  route_only_kernel<<<(a.q1 - a.q0 + kWarpsPerBlock - 1) / kWarpsPerBlock, kWarpsPerBlock * 32, 0, st>>>(a);
  return 1;
}

// ---------------------------------------------------------------------------------------
// Peer exchange: write this rank's candidate lists straight into every rank's exchange buffer
// (NVLink stores through IPC-mapped pointers), then raise this rank's flag on each of them.
// ---------------------------------------------------------------------------------------
struct PublishArgs {
  int m, world;
  const qre_candidate_t* cand;
  qre_candidate_t* dst[16];      // per peer: start of [slot][my rank] in that peer's buffer
  uint32_t* flag[16];            // per peer: my flag for this slot in that peer's buffer
  uint32_t epoch;
  unsigned* ticket;
  int per;                       // sliced results: peer p only needs queries [p * per, (p + 1) * per); 0 = all
};

__global__ void __launch_bounds__(256) publish_kernel(const PublishArgs a) {
  __shared__ unsigned s_ticket;
  const long long n16 = static_cast<long long>(a.m) * kTopK * 2;      // int4 words per list
  const int4* src = reinterpret_cast<const int4*>(a.cand);
  for (int p = 0; p < a.world; ++p) {
    int4* d = reinterpret_cast<int4*>(a.dst[p]);
    long long i0 = 0, i1 = n16;
    if (a.per > 0) {               // all-to-all of slices instead of an all-gather of whole lists
      const long long q0 = static_cast<long long>(p) * a.per, q1 = q0 + a.per;
      i0 = (q0 < a.m ? q0 : a.m) * kTopK * 2;
      i1 = (q1 < a.m ? q1 : a.m) * kTopK * 2;
    }
    for (long long i = i0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < i1;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
      d[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) s_ticket = atomicAdd(a.ticket, 1u);
  __syncthreads();
  if (s_ticket != gridDim.x - 1) return;
  __threadfence_system();                      // every CTA's stores are ordered before the flags
  if (threadIdx.x < a.world) *reinterpret_cast<volatile uint32_t*>(a.flag[threadIdx.x]) = a.epoch;
  if (threadIdx.x == 0) *a.ticket = 0u;
}

int launch_publish(int m, int world, const qre_candidate_t* cand, qre_candidate_t* const* dst, uint32_t* const* flag,
                   uint32_t epoch, unsigned* ticket, cudaStream_t st, int per) {
  PublishArgs a;
  a.m = m; a.world = world; a.cand = cand; a.epoch = epoch; a.ticket = ticket; a.per = per;
  for (int p = 0; p < world; ++p) { a.dst[p] = dst[p]; a.flag[p] = flag[p]; }
  publish_kernel<<<16, 256, 0, st>>>(a);
  return 1;
}

}  // namespace qre
