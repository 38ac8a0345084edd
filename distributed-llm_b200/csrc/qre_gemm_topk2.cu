// qre_gemm_topk2.cu — CTA-pair (cta_group::2) variant of the scan kernel.
//
// Same contract as qre_gemm_topk_kernel (qre_gemm_topk.cu): approx[q, row] on the tensor cores,
// fused per-query top-k, results in the `partial` scratch.  What changes is the data movement:
// a cluster of two CTAs (one TPC) works on 256 queries x 256 cache rows per tile with ONE
// tcgen05.mma.cta_group::2 (M = 256), each CTA keeping its own 128-query A block resident and
// loading only HALF of every cache-row tile (128 rows x 64 k = 16 KB per stage).  Per SM that
// halves the L2->SM traffic and the shared-memory reads of the B operand — the resources the
// single-CTA kernel saturates (profiles/r1_summary.md: 12.6 TB/s L2->SM) — and leaves room for
// an 8-deep B ring.
//
// Protocol (barrier objects live at the same shared-memory offset in both CTAs):
//   full[s], a_full[kb]   leader (rank 0) only, count 2: both CTAs' producers arrive (the
//                     leader with expect_tx for both halves); both CTAs' TMA loads complete_tx there.
//   empty[s], a_empty[kb], t_full[acc]   in each CTA, signalled by the leader's MMA thread with
//                     a multicast tcgen05.commit (mask 0b11).
// The resident query block is handed over per k-block: a_empty[kb] is committed right after the
// LAST tile's MMAs on k-block kb, so the next work unit's A[kb] streams in (interleaved with the
// first tile's cache-row loads) while the last tile is still being multiplied — no reload bubble
// between units.
//   t_empty[acc]      leader only, count 256: the epilogue threads of both CTAs arrive remotely.
// Only the leader issues MMAs; both CTAs run a TMA producer and four epilogue warps over their
// own 128 TMEM lanes.
#include <stdio.h>

#include <algorithm>
#include <cmath>
#include <type_traits>

#include "qre_common.cuh"
#include "qre_encoder.cuh"      // programmatic dependent launch helpers
#include "qre_ptx.cuh"
#include "qre_scan_epilogue.cuh"

namespace qre {

namespace {

constexpr int kThreads2 = 192;
constexpr int kStages2 = 7;           // 7 x 16 KB: leaves ~18 KB of the SM's shared memory to small kernels running beside the scan
constexpr int kABytesPerKB2 = kBM * kBK * 2;          // 16 KB
constexpr int kBHalfRows = kBN / 2;                   // 128 cache rows per CTA per tile
constexpr int kBHalfBytes = kBHalfRows * kBK * 2;     // 16 KB
constexpr int kSmemA2 = kMaxKB * kABytesPerKB2;       // 96 KB
constexpr int kSmemB2 = kStages2 * kBHalfBytes;       // 128 KB
constexpr int kSmemTotal2 = kSmemA2 + kSmemB2 + 512 + 1024;
constexpr uint32_t kTmemCols2 = 512;
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;           // shared::cluster address of the even CTA of a pair

// ---- cluster / cta_group::2 PTX ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive (count 1, no tx) on the barrier at the same offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t rank) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
      :
      : "r"(smem_u32(bar)), "r"(rank)
      : "memory");
}
// TMA load into THIS CTA's shared memory, completion bytes credited to the pair leader's barrier
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int32_t c0,
                                                 int32_t c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      :
      : "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerMask), "r"(c0),
        "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish2() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_ss_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                  uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :
      : "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs once all MMAs issued so far have retired
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      :
      : "r"(smem_u32(bar)), "h"(static_cast<uint16_t>(3))
      : "memory");
}


// ---- work-unit queue (dynamic scheduling) ----
// The pair's leader producer claims work units from a global counter and hands each id to every
// role of both CTAs through a small ring in each CTA's shared memory: value store + release
// arrive at cluster scope, the readers wait with acquire at cluster scope.  The ring is never
// overrun without an "empty" barrier: a reader lags the claimer by at most the B ring (7 tiles),
// two accumulators and two units in flight, and a unit is at least one tile (< kUQ).
constexpr int kUQ = 16;
__device__ __forceinline__ void uq_publish(uint64_t* bar, int* val, uint32_t k, int u) {
  const uint32_t b = smem_u32(&bar[k % kUQ]), v = smem_u32(&val[k % kUQ]);
#pragma unroll
  for (uint32_t r = 0; r < 2; ++r)
    asm volatile(
        "{\n\t.reg .b32 ra, rb;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %2;\n\t"
        "mapa.shared::cluster.u32 rb, %1, %2;\n\t"
        "st.shared::cluster.u32 [ra], %3;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [rb];\n\t}"
        :
        : "r"(v), "r"(b), "r"(r), "r"(u)
        : "memory");
}
__device__ __forceinline__ int uq_take(uint64_t* bar, const int* val, uint32_t k) {
  const uint32_t b = smem_u32(&bar[k % kUQ]), parity = (k / kUQ) & 1u;
  uint32_t spins = 0;
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(b), "r"(parity)
        : "memory");
    if (ok) break;
    if (++spins > (1u << 26)) {
      printf("qre: work-unit queue wait timed out (block %d thread %d k %u)\n", blockIdx.x, threadIdx.x, k);
      __trap();
    }
  }
  return *reinterpret_cast<const volatile int*>(&val[k % kUQ]);
}

}  // namespace

template <bool kDump>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
qre_gemm_topk_pair_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_eh,
                          const __grid_constant__ GemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kSmemA2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSmemA2 + kSmemB2);
  uint64_t* full = bars;                       // [kStages2]
  uint64_t* empty = bars + kStages2;           // [kStages2]
  uint64_t* a_full = bars + 2 * kStages2;      // [kMaxKB]
  uint64_t* a_empty = a_full + kMaxKB;         // [kMaxKB]
  uint64_t* t_full = a_empty + kMaxKB;         // [2]
  uint64_t* t_empty = t_full + 2;              // [2]
  uint64_t* uq_full = t_empty + 2;              // [kUQ]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(uq_full + kUQ);
  int* uq_val = reinterpret_cast<int*>(tmem_slot + 2);     // [kUQ]

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int n_clusters = gridDim.x >> 1;

  if (warp == 4 && lane == 0) {
    tma_prefetch_desc(&tmap_q);
    tma_prefetch_desc(&tmap_eh);
    for (int i = 0; i < kStages2; ++i) {
      mbar_init(&full[i], 2);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < kMaxKB; ++i) {
      mbar_init(&a_full[i], 2);
      mbar_init(&a_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], 256);
    }
    for (int i = 0; i < kUQ; ++i) mbar_init(&uq_full[i], 1);
    fence_barrier_init();
  }
  cluster_sync_all();                          // barriers of both CTAs are initialised
  if (warp == 5) {
    tmem_alloc2(tmem_slot, kTmemCols2);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                                  // set-up done; the queries (prep kernel) must be complete from here

  const int npairs = (a.mblocks + 1) >> 1;
  const int U = a.splits * npairs;
  const long long T = a.tiles_n;

  if (warp == 4) {
    // ================= TMA producer (one per CTA) =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0, a_iter = 0, k = 0;
      int u = cluster_id;                      // the first unit is static, the rest are claimed
      while (u < U) {
        const int s = u / npairs, pr = u - s * npairs;
        const int mb = 2 * pr + static_cast<int>(rank);
        const int t0 = static_cast<int>((T * s) / a.splits), t1 = static_cast<int>((T * (s + 1)) / a.splits);
        // The leader claims the pair's next unit a few tiles before this one ends (late, so that
        // pairs which finish together sweep the same row range together; early enough for the
        // atomic's latency to hide under one tile's loads and for the peer to see the id in time).
        const int t_claim = t1 - 3 > t0 ? t1 - 3 : t0;
        int u_next = U;
        for (int t = t0; t < t1; ++t) {
          unsigned claimed = 0;
          if (leader && t == t_claim) claimed = atomicAdd(a.work_ctr, 1u);
          for (int kb = 0; kb < a.num_kb; ++kb) {
            if (t == t0) {                     // this unit's query block, one k-block at a time
              mbar_wait(&a_empty[kb], (a_iter & 1) ^ 1);
              if (leader) mbar_arrive_expect_tx(&a_full[kb], 2u * kABytesPerKB2);
              else mbar_arrive_remote(&a_full[kb], 0);
              tma_load_2d_pair(sA + kb * kABytesPerKB2, &tmap_q, &a_full[kb], kb * kBK, mb * kBM);
            }
            mbar_wait(&empty[stage], phase ^ 1);
            if (leader) mbar_arrive_expect_tx(&full[stage], 2u * kBHalfBytes);
            else mbar_arrive_remote(&full[stage], 0);
            tma_load_2d_pair(sB + stage * kBHalfBytes, &tmap_eh, &full[stage], kb * kBK,
                             t * kBN + static_cast<int>(rank) * kBHalfRows);
            if (++stage == kStages2) { stage = 0; phase ^= 1; }
          }
          if (leader && t == t_claim) {
            const unsigned long long nx = static_cast<unsigned long long>(n_clusters) + claimed;
            u_next = nx < static_cast<unsigned long long>(U) ? static_cast<int>(nx) : U;
            uq_publish(uq_full, uq_val, k + 1, u_next);
          }
        }
        ++a_iter;
        ++k;
        u = leader ? u_next : uq_take(uq_full, uq_val, k);
      }
      if (leader) {
        // every pair claims exactly once past the end; the last pair to do so re-arms the counters
        // for the next launch on this batch slot
        __threadfence();
        if (atomicAdd(a.work_ctr + 1, 1u) == static_cast<unsigned>(n_clusters) - 1u) {
          a.work_ctr[0] = 0u;
          a.work_ctr[1] = 0u;
          __threadfence();
        }
      }
    }
  } else if (warp == 5) {
    // ================= MMA issuer: warp 5 of the leader CTA =================
    // The whole warp runs the loop convergently and one elected lane issues (the CUTLASS /
    // DeepGEMM shape): with warp-uniform control flow the descriptors stay in uniform registers
    // and every tcgen05.mma is one predicated instruction.  Issued from a lone lane inside a
    // divergent branch instead, the compiler wraps each one in an elect / broadcast loop, and the
    // issuing thread — not the tensor pipe — sets the pace (measured: 71 % vs 82 % pipe-active).
    // The tile loop is peeled so that only a unit's first tile waits for the query k-blocks and
    // only its last tile hands them back; the steady-state tile has no conditionals at all.
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * kBM, kBN);      // M = 256 across the pair
      uint32_t stage = 0, phase = 0, a_iter = 0, tile_iter = 0;
      const uint32_t sA_addr = smem_u32(sA), sB_addr = smem_u32(sB);
      const int num_kb = a.num_kb;
      auto issue_tile = [&](auto first_c, auto last_c) {
        constexpr bool kFirst = decltype(first_c)::value, kLast = decltype(last_c)::value;
        const uint32_t acc = tile_iter & 1;
        mbar_wait(&t_empty[acc], ((tile_iter >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kBN;
        for (int kb = 0; kb < num_kb; ++kb) {
          if (kFirst) mbar_wait(&a_full[kb], a_iter & 1);
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t adesc = umma_desc_sw128(sA_addr + kb * kABytesPerKB2);
            const uint64_t bdesc = umma_desc_sw128(sB_addr + stage * kBHalfBytes);
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k)
              umma_bf16_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit_pair(&empty[stage]);
            if (kLast) umma_commit_pair(&a_empty[kb]);               // last reader of A[kb] in this unit
            if (kb == num_kb - 1) umma_commit_pair(&t_full[acc]);
          }
          __syncwarp();
          if (++stage == kStages2) { stage = 0; phase ^= 1; }
        }
        ++tile_iter;
      };
      using T_ = std::true_type;
      using F_ = std::false_type;
      uint32_t k = 0;
      int u = cluster_id;
      while (u < U) {
        const int s = u / npairs;
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
        u = uq_take(uq_full, uq_val, k);
      }
    }
  } else {
    // ================= epilogue (both CTAs): fused per-query top-k =================
    uint32_t tile_iter = 0, k = 0;
    TopK tk;
    Cursor cu;
    int u = cluster_id;
    while (u < U) {
      const int s = u / npairs, pr = u - s * npairs;
      const int mb = 2 * pr + static_cast<int>(rank);
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
        mbar_arrive_remote(&t_empty[acc], 0);      // the leader's MMA thread owns the accumulator hand-off
        ++tile_iter;
      }
      if (active) topk_store(cu, tk, a, static_cast<long long>(q) * a.splits + s);
      ++k;
      u = uq_take(uq_full, uq_val, k);
    }
  }

  __syncwarp();                                // the single-lane roles rejoin their warps
  tc_fence_before();
  cluster_sync_all();                          // nobody frees TMEM / exits while the peer still uses it
  if (warp == 5) tmem_dealloc2(tmem_base, kTmemCols2);
}

cudaError_t gemm_topk_pair_init() {
  cudaError_t e = cudaFuncSetAttribute(qre_gemm_topk_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       kSmemTotal2);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(qre_gemm_topk_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              kSmemTotal2);
}

// Same search as plan_scan, over query-block PAIRS and CTA-pair clusters.
ScanGeometry plan_scan_pair(int m, long long n_rows, int num_sms) {
  ScanGeometry g = plan_scan(m, n_rows, num_sms);
  const int npairs = (g.mblocks + 1) / 2;
  const int clusters = num_sms / 2;
  long long S = 1;
  double best = -1.0;
  const long long s_max = std::min<long long>(g.tiles_n, (8LL * clusters) / npairs + 1);
  for (long long c = 1; c <= s_max; ++c) {
    const double U = static_cast<double>(c) * npairs;
    const double rounds = std::ceil(U / clusters);
    if (rounds > 8.0) break;
    const double L = static_cast<double>(g.tiles_n) / c;
    const double score = (U / clusters) / rounds * L / (L + 1.6);
    if (score > best + 1e-12) { best = score; S = c; }
  }
  g.splits = static_cast<int>(S);
  const long long U = S * npairs;
  g.grid = 2 * static_cast<int>(U < clusters ? U : clusters);
  return g;
}

int launch_gemm_topk_pair(DeviceState& s, int m, long long n_rows, const qre_params_t& p,
                          float* dump, long long dump_ld, long long dump_row0, cudaStream_t st) {
  ScanGeometry g = plan_scan_pair(m, n_rows, scan_sms(s, p));
  g.pair = true;
  const GemmArgs a = make_gemm_args(s, g, m, n_rows, p, dump, dump_ld, dump_row0);
  const bool prof = s.prof_on && s.prof_ev != nullptr && s.prof_used < kProfMax && dump == nullptr;
  if (prof) cudaEventRecord(s.prof_ev[2 * s.prof_used], st);
  if (dump != nullptr)
    launch_chain(qre_gemm_topk_pair_kernel<true>, dim3(g.grid), dim3(kThreads2), kSmemTotal2, st, s.pdl,
                 s.cur().tmap_q, s.tmap_eh, a);
  else
    launch_chain(qre_gemm_topk_pair_kernel<false>, dim3(g.grid), dim3(kThreads2), kSmemTotal2, st, s.pdl,
                 s.cur().tmap_q, s.tmap_eh, a);
  if (prof) {
    cudaEventRecord(s.prof_ev[2 * s.prof_used + 1], st);
    ++s.prof_used;
  }
  return 1;
}

}  // namespace qre
