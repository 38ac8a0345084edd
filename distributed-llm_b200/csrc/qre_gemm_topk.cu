// qre_gemm_topk.cu — the dominant kernel of the routing-cache lookup.
//
// Replaces the reference's per-entry cosine loop (src/cache.py:287-293) for a whole batch:
// approx[q, row] = <bf16(q/|q|), bf16(e/|e|)> on the 5th-gen tensor cores, fused with a
// per-query running top-k so the [m x N] score matrix never leaves the SM.
//
//   * one CTA per SM, persistent over work units (row range s, query block mb);
//   * for dim <= 384 the 128-query A block (128 x dim bf16 <= 96 KB) is loaded ONCE per unit by
//     TMA and stays resident in shared memory (for larger dims A k-blocks ride with B); only cache-row tiles (256 rows x 64 k, 32 KB) stream
//     through a 4-stage TMA/mbarrier ring;
//   * a single elected thread issues tcgen05.mma (M=128, N=256, K=16, bf16 -> fp32) into one
//     of two 256-column TMEM accumulators, so the epilogue of tile t overlaps the MMAs of t+1;
//   * four epilogue warps read their 32 TMEM lanes with tcgen05.ld (one lane = one query),
//     reduce each 64-column chunk to its max and compare with the query's gate (initialised
//     to the hit floor thr-eps): only chunks that can still matter take the slow path (append
//     the candidate to the (query, range)'s list; the context / TTL mask from RowMeta is applied
//     by the re-score kernel, qre_scan_epilogue.cuh).
//
// Units are ordered row-range-major (u = s * mblocks + mb); a CTA's first unit is static, the
// rest are claimed from a global counter, so at any moment the resident CTAs sweep the SAME rows
// for different query blocks — also when some units take longer than others — and the 32
// re-reads of each cache-row tile are served by L2, not HBM.
#include <stdio.h>

#include <algorithm>
#include <cmath>
#include <type_traits>

#include "qre_common.cuh"
#include "qre_encoder.cuh"      // programmatic dependent launch helpers
#include "qre_ptx.cuh"
#include "qre_scan_epilogue.cuh"

namespace qre {

constexpr int kThreads = 192;                       // warps 0-3 epilogue, 4 TMA, 5 MMA
constexpr int kABytesPerKB = kBM * kBK * 2;         // 16 KB
constexpr int kBBytes = kBN * kBK * 2;              // 32 KB
constexpr int kSmemA = kMaxKB * kABytesPerKB;       // 96 KB
constexpr int kSmemB = kStages * kBBytes;           // 128 KB
constexpr int kSmemBar = 512;
constexpr int kSmemTotal = kSmemA + kSmemB + kSmemBar + 1024;   // + alignment slack
constexpr uint32_t kTmemCols = 512;

// Work units are claimed dynamically (see qre_gemm_topk2.cu): the producer thread takes the next
// id from a global counter and hands it to the MMA and epilogue warps through a ring in shared
// memory.  A reader lags the claimer by at most the B ring, two accumulators and two units in flight.
constexpr int kUQ1 = 16;
__device__ __forceinline__ void uq1_publish(uint64_t* bar, int* val, uint32_t k, int u) {
  *reinterpret_cast<volatile int*>(&val[k % kUQ1]) = u;
  mbar_arrive(&bar[k % kUQ1]);                      // release at CTA scope orders the store above
}
__device__ __forceinline__ int uq1_take(uint64_t* bar, const int* val, uint32_t k) {
  mbar_wait(&bar[k % kUQ1], (k / kUQ1) & 1u);
  return *reinterpret_cast<const volatile int*>(&val[k % kUQ1]);
}

// kResident: the 128-query A block fits in shared memory next to the B ring (dim_pad <= 384)
// and is loaded once per work unit.  Otherwise (dim_pad up to 1024, e.g. 768-d embeddings) each
// pipeline stage carries the A k-block (16 KB) together with the B k-block (32 KB).
template <bool kDump, bool kResident>
__global__ void __launch_bounds__(kThreads, 1)
qre_gemm_topk_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_e,
                     const __grid_constant__ GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  // resident: [A: 6 x 16 KB][B ring: 4 x 32 KB]; streaming: ring of 4 x (A 16 KB + B 32 KB)
  constexpr int kStageBytes = kResident ? kBBytes : (kABytesPerKB + kBBytes);
  uint8_t* sA = smem;
  uint8_t* sB = kResident ? smem + kSmemA : smem + kABytesPerKB;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (kResident ? kSmemA + kSmemB : kStages * kStageBytes));
  uint64_t* full = bars;                    // [kStages]  TMA -> MMA
  uint64_t* empty = bars + kStages;         // [kStages]  MMA -> TMA
  uint64_t* a_full = bars + 2 * kStages;    // [kMaxKB] TMA -> MMA (resident A k-block landed)
  uint64_t* a_empty = a_full + kMaxKB;      // [kMaxKB] MMA -> TMA (last MMA of the unit reading A[kb] retired)
  uint64_t* t_full = a_empty + kMaxKB;      // [2] MMA -> epilogue
  uint64_t* t_empty = t_full + 2;           // [2] epilogue -> MMA
  uint64_t* uq_full = t_empty + 2;          // [kUQ1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(uq_full + kUQ1);
  int* uq_val = reinterpret_cast<int*>(tmem_slot + 2);     // [kUQ1]

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_e);
    for (int i = 0; i < kUQ1; ++i) mbar_init(&uq_full[i], 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kMaxKB; ++i) {
      mbar_init(&a_full[i], 1);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 5) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                               // set-up done; the queries (prep kernel) must be complete from here

  const int U = a.splits * a.mblocks;
  const long long T = a.tiles_n;

  if (warp == 4) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0, a_iter = 0, k = 0;
      int u = blockIdx.x;                        // the first unit is static, the rest are claimed
      while (u < U) {
        const int s = u / a.mblocks, mb = u - s * a.mblocks;
        const int t0 = static_cast<int>((T * s) / a.splits), t1 = static_cast<int>((T * (s + 1)) / a.splits);
        const int t_claim = t1 - 3 > t0 ? t1 - 3 : t0;
        int u_next = U;
        for (int t = t0; t < t1; ++t) {
          unsigned claimed = 0;
          if (t == t_claim) claimed = atomicAdd(a.work_ctr, 1u);
          for (int kb = 0; kb < a.num_kb; ++kb) {
            if (kResident && t == t0) {        // this unit's query block, handed over per k-block
              mbar_wait(&a_empty[kb], (a_iter & 1) ^ 1);
              mbar_arrive_expect_tx(&a_full[kb], kABytesPerKB);
              tma_load_2d(sA + kb * kABytesPerKB, &tmap_q, &a_full[kb], kb * kBK, mb * kBM);
            }
            mbar_wait(&empty[stage], phase ^ 1);
            mbar_arrive_expect_tx(&full[stage], kStageBytes);
            if (!kResident) tma_load_2d(sA + stage * kStageBytes, &tmap_q, &full[stage], kb * kBK, mb * kBM);
            tma_load_2d(sB + stage * kStageBytes, &tmap_e, &full[stage], kb * kBK, t * kBN);
            if (++stage == kStages) { stage = 0; phase ^= 1; }
          }
          if (t == t_claim) {
            const unsigned long long nx = static_cast<unsigned long long>(gridDim.x) + claimed;
            u_next = nx < static_cast<unsigned long long>(U) ? static_cast<int>(nx) : U;
            uq1_publish(uq_full, uq_val, k + 1, u_next);
          }
        }
        ++a_iter;
        ++k;
        u = u_next;
      }
      // every CTA claims exactly once past the end; the last one re-arms the counters
      __threadfence();
      if (atomicAdd(a.work_ctr + 1, 1u) == gridDim.x - 1u) {
        a.work_ctr[0] = 0u;
        a.work_ctr[1] = 0u;
        __threadfence();
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer =================
    // Whole warp in the loop, one elected lane issues (see qre_gemm_topk2.cu for why); the tile loop
    // is peeled so only a unit's first / last tile touch the query-block hand-off barriers.
    {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, kBN);
      uint32_t stage = 0, phase = 0, a_iter = 0, tile_iter = 0;
      const uint32_t sA_addr = smem_u32(sA), sB_addr = smem_u32(sB);
      const int num_kb = a.num_kb;
      auto issue_tile = [&](auto first_c, auto last_c) {
        constexpr bool kFirst = kResident && decltype(first_c)::value;
        constexpr bool kLast = kResident && decltype(last_c)::value;
        const uint32_t acc = tile_iter & 1;
        mbar_wait(&t_empty[acc], ((tile_iter >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kBN;
        for (int kb = 0; kb < num_kb; ++kb) {
          if (kFirst) mbar_wait(&a_full[kb], a_iter & 1);
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t adesc = umma_desc_sw128(sA_addr + (kResident ? kb * kABytesPerKB : stage * kStageBytes));
            const uint64_t bdesc = umma_desc_sw128(sB_addr + stage * kStageBytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k) {
              // advance 16 bf16 = 32 bytes along K inside the 128-byte swizzle row: +2 in (addr>>4)
              umma_bf16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit(&empty[stage]);                    // stage free once these MMAs retire
            if (kLast) umma_commit(&a_empty[kb]);          // last reader of A[kb] in this unit
            if (kb == num_kb - 1) umma_commit(&t_full[acc]);   // accumulator complete -> epilogue
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        ++tile_iter;
      };
      using T_ = std::true_type;
      using F_ = std::false_type;
      uint32_t k = 0;
      int u = blockIdx.x;
      while (u < U) {
        const int s = u / a.mblocks;
        const int t0 = static_cast<int>((T * s) / a.splits), t1 = static_cast<int>((T * (s + 1)) / a.splits);
        if (t1 - t0 == 1) {
          issue_tile(T_{}, T_{});
        } else {
          issue_tile(T_{}, F_{});
          for (int t = t0 + 1; t < t1 - 1; ++t) issue_tile(F_{}, F_{});
          issue_tile(F_{}, T_{});
        }
        ++a_iter;
        ++k;
        u = uq1_take(uq_full, uq_val, k);
      }
    }
  } else {
    // ================= epilogue: fused per-query top-k =================
    uint32_t tile_iter = 0, k = 0;
    TopK tk;
    Cursor cu;
    int u = blockIdx.x;
    while (u < U) {
      const int s = u / a.mblocks, mb = u - s * a.mblocks;
      const int t0 = static_cast<int>((T * s) / a.splits), t1 = static_cast<int>((T * (s + 1)) / a.splits);
      const int q = mb * kBM + warp * 32 + lane;
      const bool active = q < a.m;
      const uint64_t myctx = active ? a.q_ctx[q] : 0ull;
      const float kth = active ? a.floor_v : __int_as_float(0x7f800000);
      topk_reset(cu, tk, a, static_cast<long long>(q) * a.splits + s, kth);
      for (int t = t0; t < t1; ++t) {
        const uint32_t acc = tile_iter & 1;
        mbar_wait(&t_full[acc], (tile_iter >> 1) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + acc * kBN;
        scan_tile<kDump>(taddr, static_cast<long long>(t) * kBN, q, active, cu, tk, myctx, a);
        tc_fence_before();
        mbar_arrive(&t_empty[acc]);
        ++tile_iter;
      }
      if (active) topk_store(cu, tk, a, static_cast<long long>(q) * a.splits + s);
      ++k;
      u = uq1_take(uq_full, uq_val, k);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 5) tmem_dealloc(tmem_base, kTmemCols);
}

// Work decomposition: S row ranges per M-block so that S*mblocks is about `waves` full waves
// of the SM count, with units of >= 16 tiles where the problem allows it.
ScanGeometry plan_scan(int m, long long n_rows, int num_sms) {
  ScanGeometry g;
  g.mblocks = (m + kBM - 1) / kBM;
  g.tiles_n = static_cast<int>((n_rows + kBN - 1) / kBN);
  if (g.mblocks < 1) g.mblocks = 1;
  if (g.tiles_n < 1) g.tiles_n = 1;
  // Units are equal-sized, so a launch takes ceil(U / G) rounds of T / S tiles; every unit also
  // pays ~1.6 tile-times of pipeline fill (A load, first B stages).  Pick the S that maximises
  //   (U / G) / ceil(U / G)  *  L / (L + 1.6),   L = T / S,  U = S * mblocks,
  // among S giving at most 8 rounds.
  long long S = 1;
  double best = -1.0;
  const long long s_max = std::min<long long>(g.tiles_n, (8LL * num_sms) / g.mblocks + 1);
  for (long long c = 1; c <= s_max; ++c) {
    const double U = static_cast<double>(c) * g.mblocks;
    const double rounds = std::ceil(U / num_sms);
    if (rounds > 8.0) break;
    const double L = static_cast<double>(g.tiles_n) / c;
    const double score = (U / num_sms) / rounds * L / (L + 1.6);
    if (score > best + 1e-12) { best = score; S = c; }
  }
  g.splits = static_cast<int>(S);
  const long long U = S * g.mblocks;
  g.grid = static_cast<int>(U < num_sms ? U : num_sms);
  return g;
}

// SMs the scan may occupy: all of them minus what the caller reserves for kernels that should
// run beside it (NCCL all-gather + merge of the previous batch in the pipelined multi-GPU path).
int scan_sms(const DeviceState& s, const qre_params_t& p) {
  int r = p.reserve_sms;
  if (r < 0) r = 0;
  if (r > s.num_sms / 2) r = s.num_sms / 2;
  return s.num_sms - r;
}

ScanGeometry plan_for(const DeviceState& s, int m, long long n_rows, const qre_params_t& p) {
  const int sms = scan_sms(s, p);
  const int mblocks = (m + kBM - 1) / kBM;
  if (s.allow_pair && s.dim_pad / kBK <= kMaxKB && mblocks >= 2 && sms >= 2) {
    ScanGeometry g = plan_scan_pair(m, n_rows, sms);
    g.pair = true;
    return g;
  }
  return plan_scan(m, n_rows, sms);
}

// Per-device set-up (called by qre_cache_create with the device current): opt in to the
// 225 KB dynamic shared memory the scan kernel needs.
template <bool kDump, bool kResident>
static cudaError_t set_smem_attr() {
  return cudaFuncSetAttribute(qre_gemm_topk_kernel<kDump, kResident>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              kSmemTotal);
}
cudaError_t gemm_topk_init() {
  cudaError_t e = set_smem_attr<true, true>();
  if (e == cudaSuccess) e = set_smem_attr<false, true>();
  if (e == cudaSuccess) e = set_smem_attr<true, false>();
  if (e == cudaSuccess) e = set_smem_attr<false, false>();
  return e;
}

GemmArgs make_gemm_args(const DeviceState& s, const ScanGeometry& g, int m, long long n_rows, const qre_params_t& p,
                        float* dump, long long dump_ld, long long dump_row0) {
  const DeviceState::Batch& b = s.cur();
  GemmArgs a;
  a.m = m;
  a.mblocks = g.mblocks;
  a.tiles_n = g.tiles_n;
  a.splits = g.splits;
  a.num_kb = s.dim_pad / kBK;
  a.n_rows = n_rows;
  const double thr = p.similarity_threshold > 0.0 ? p.similarity_threshold : 0.0;
  a.floor_v = static_cast<float>(thr) - kApproxEps;
  a.now_us = p.now_us;
  a.ttl_us = p.ttl_us;
  a.q_ctx = b.q_ctx;
  a.meta = s.meta;
  a.partial = b.partial;
  a.spill_head = b.spill_head;
  a.spill_pool = b.spill_pool;
  a.spill_next = b.counters;
  a.spill_cap = s.spill_cap;
  a.work_ctr = b.counters + 4;
  a.dump = dump;
  a.dump_ld = dump_ld;
  a.dump_row0 = dump_row0;
  return a;
}

// The scratch is sized from the hard bound on work units (8 rounds of the SM count); a geometry
// that would not fit is a bug in the planner, never a silent overrun.
bool geometry_fits(const DeviceState& s, const ScanGeometry& g) {
  const long long blocks = g.pair ? 2LL * ((g.mblocks + 1) / 2) : g.mblocks;
  return blocks * kBM * g.splits * kTopK <= s.partial_entries;
}

int launch_gemm_topk(DeviceState& s, int m, long long n_rows, const qre_params_t& p,
                     float* dump, long long dump_ld, long long dump_row0, int dump_n, cudaStream_t st) {
  (void)dump_n;
  const ScanGeometry g = plan_for(s, m, n_rows, p);
  if (!geometry_fits(s, g)) return -1;
  if (g.pair) return launch_gemm_topk_pair(s, m, n_rows, p, dump, dump_ld, dump_row0, st);
  const GemmArgs a = make_gemm_args(s, g, m, n_rows, p, dump, dump_ld, dump_row0);
  const bool prof = s.prof_on && s.prof_ev != nullptr && s.prof_used < kProfMax && dump == nullptr;
  if (prof) cudaEventRecord(s.prof_ev[2 * s.prof_used], st);
  const bool resident = a.num_kb <= kMaxKB;
  if (dump != nullptr) {
    if (resident) launch_chain(qre_gemm_topk_kernel<true, true>, dim3(g.grid), dim3(kThreads), kSmemTotal, st, s.pdl, s.cur().tmap_q, s.tmap_e, a);
    else launch_chain(qre_gemm_topk_kernel<true, false>, dim3(g.grid), dim3(kThreads), kSmemTotal, st, s.pdl, s.cur().tmap_q, s.tmap_e, a);
  } else {
    if (resident) launch_chain(qre_gemm_topk_kernel<false, true>, dim3(g.grid), dim3(kThreads), kSmemTotal, st, s.pdl, s.cur().tmap_q, s.tmap_e, a);
    else launch_chain(qre_gemm_topk_kernel<false, false>, dim3(g.grid), dim3(kThreads), kSmemTotal, st, s.pdl, s.cur().tmap_q, s.tmap_e, a);
  }
  if (prof) {
    cudaEventRecord(s.prof_ev[2 * s.prof_used + 1], st);
    ++s.prof_used;
  }
  return 1;
}

}  // namespace qre
