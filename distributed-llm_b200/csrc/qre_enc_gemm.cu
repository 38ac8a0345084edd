// qre_enc_gemm.cu — the linear layers of the on-GPU query embedder:
//   out[token, f] = epilogue( sum_k x[token, k] * W[f, k] + bias[f] )        bf16 x bf16 -> fp32
// for the four projections of a BERT layer (QKV, attention output, FFN up, FFN down), which the
// reference reaches through SentenceTransformer.encode (src/query_router_engine.py:181, 571).
//
//   * tokens are the M operand, one TMEM lane per token; a cluster of two CTAs (one TPC) works on
//     256 tokens x 192 features per tile with ONE tcgen05.mma.cta_group::2 (M = 256): each CTA
//     streams its own 128 tokens of x and only HALF of the weight tile (96 rows).  The single-CTA
//     form of this kernel ran at the L2->SM limit (ncu: 1.1 GB of L2 reads per QKV projection,
//     11.5 TB/s), not at the tensor pipe's; pairing removes 30 % of that traffic;
//   * W is in torch.nn.Linear layout [out, in] = K-major, exactly what the UMMA B operand wants:
//     no transposition anywhere.  192 output features per tile (see qre_encoder.cuh);
//   * persistent clusters; a 6-stage TMA/mbarrier ring carries 128x64 of x and 96x64 of W per
//     CTA per stage (28 KB); one elected thread of the leader CTA issues the MMAs into one of two
//     TMEM accumulators, so the epilogue of tile i runs under the MMAs of tile i+1;
//   * twelve epilogue warps per CTA, one per (TMEM lane quarter, 64-column group).  Each owns a
//     4 KB staging box in shared memory (32 tokens x 64 features, 128-byte swizzle): the residual
//     rows arrive in it by TMA while the MMAs run, the warp adds accumulator + bias (+ GELU |
//     + residual) in place, and one TMA store writes the box out — every byte of global traffic is
//     a full line moved by the copy engine.
//
// Barrier protocol (objects at the same shared-memory offset in both CTAs), as in
// qre_gemm_topk2.cu: full[s] on the leader (count 2: both producers arrive, the leader with the
// expect_tx of both CTAs' bytes); empty[s] and t_full[acc] in each CTA, signalled by a multicast
// tcgen05.commit; t_empty[acc] on the leader (count 768: the epilogue threads of both CTAs).
//
// Tiles are walked feature-tile-fastest, so the clusters resident at any moment share a few token
// blocks (L2 hits on x) and all of W (<= 1.2 MB, L2-resident).
#include <stdio.h>

#include "qre_encoder.cuh"
#include "qre_ptx.cuh"

namespace qre {

namespace {

constexpr int kEpiWarps = 12;                          // warps 0-11 epilogue, 12 TMA, 13 MMA
constexpr int kThreads = (kEpiWarps + 2) * 32;
constexpr int kABytes = kEncBM * kEncBK * 2;           // 16 KB: this CTA's 128 tokens
constexpr int kWHalfRows = kEncBN / 2;                 // 96 weight rows per CTA per tile
constexpr int kBBytes = kWHalfRows * kEncBK * 2;       // 12 KB
constexpr int kStageBytes = kABytes + kBBytes;         // 28 KB per CTA
constexpr int kBoxBytes = 32 * 64 * 2;                 // 4 KB staging box per epilogue warp
constexpr int kSmemBar = 512;
constexpr int kLnStages = 4;                           // the LayerNorm epilogue trades two ring stages for a second box
constexpr int kLnStatBytes = 2 * kEpiWarps * 32 * 8;   // double-buffered per-row (sum, sum of squares) partials
constexpr int smem_total(int epi) {
  return epi == kEpiResidualLN
             ? kLnStages * kStageBytes + 2 * kEpiWarps * kBoxBytes + kSmemBar + kLnStatBytes + 1024
             : kEncStages * kStageBytes + kEpiWarps * kBoxBytes + kSmemBar + 1024;
}
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kAccStride = 256;                   // TMEM column offset of the second accumulator
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;            // shared::cluster address of the even CTA of a pair

// ---- cluster / cta_group::2 PTX (same forms as qre_gemm_topk2.cu) ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
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

// GELU with the exact (erf) definition, x * Phi(x), for two values at once.  2 Phi(-a) = exp2(R(a)) for
// a = |x| in [0, 5.25], R a degree-6 polynomial fitted to 1 + log2(0.5 erfc(a / sqrt 2)) (near-minimax,
// Chebyshev nodes); then x Phi(x) = h + |h| (1 - 2 Phi(-|x|)) with h = x / 2, which needs no select.
// Measured against float64 erf over [-12, 12]: |error| <= 2.8e-6 absolute, 2.1e-5 relative wherever
// |gelu| > 1e-6 — two orders below the bf16 rounding that follows.  Per pair: 9 packed FMA/MUL, 2 MUFU
// (ex2) and 4 scalar instructions.  The FFN-up epilogue is bound by instruction issue (24,576 elements per
// tile against ~2 us of tensor-core time); erff would cost four times that.
__device__ __forceinline__ float2 gelu_erf2(float2 x) {
  const float2 a = make_float2(fminf(fabsf(x.x), 5.25f), fminf(fabsf(x.y), 5.25f));
  float2 r = ffma2(a, splat(2.802607560e-05f), splat(-6.891732337e-04f));
  r = ffma2(a, r, splat(7.647311781e-03f));
  r = ffma2(a, r, splat(-5.238423496e-02f));
  r = ffma2(a, r, splat(-4.598393440e-01f));
  r = ffma2(a, r, splat(-1.150800705e+00f));
  r = ffma2(a, r, splat(-1.943111420e-05f));               // constant term of R: 1 - 1.000019431
  const float2 e2 = make_float2(ex2_approx(r.x), ex2_approx(r.y));      // 2 Phi(-|x|); -27 < r <= 0
  const float2 t = ffma2(e2, splat(-1.0f), splat(1.0f));
  const float2 h = fmul2(x, splat(0.5f));
  return ffma2(make_float2(fabsf(h.x), fabsf(h.y)), t, h);
}

template <int kEpi>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
qre_enc_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                    const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                    const EncGemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr bool kLN = kEpi == kEpiResidualLN;
  constexpr int kStages = kLN ? kLnStages : kEncStages;
  constexpr int kBoxesPerWarp = kLN ? 2 : 1;
  uint8_t* boxes = smem + kStages * kStageBytes;                    // [kEpiWarps][kBoxesPerWarp] x 4 KB, 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(boxes + kEpiWarps * kBoxesPerWarp * kBoxBytes);
  float2* ln_stats = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(bars) + kSmemBar);   // kLN only
  uint64_t* full = bars;                        // [kStages] TMA -> MMA
  uint64_t* empty = bars + kStages;             // [kStages] MMA -> TMA
  uint64_t* t_full = bars + 2 * kStages;        // [2] MMA -> epilogue
  uint64_t* t_empty = t_full + 2;               // [2] epilogue -> MMA
  uint64_t* res_bar = t_empty + 2;              // [kEpiWarps] residual box landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kEpiWarps);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1;
  const int n_clusters = gridDim.x >> 1;

  if (warp == kEpiWarps && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_out);
    if (kEpi == kEpiResidual || kLN) tma_prefetch_desc(&tmap_res);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full[i], 2);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], 2 * kEpiWarps * 32);
    }
    for (int i = 0; i < kEpiWarps; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  cluster_sync_all();                          // barriers of both CTAs are initialised
  if (warp == kEpiWarps + 1) {
    tmem_alloc2(tmem_slot, kTmemCols);
    tmem_relinquish2();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_launch_dependents();                     // the next kernel of the chain may start its own set-up
  pdl_wait();                                  // everything above overlapped the previous kernel's tail
  const int npairs = (a.tiles_m + 1) >> 1;     // pairs of 128-token blocks
  // The j-th tile of this cluster.  Default: tiles dealt round-robin, feature tile fastest.  LayerNorm
  // epilogue: whole 256-token units (both feature tiles, so one CTA sees complete rows), units round-robin.
  const int n_mine = kLN ? (cluster_id < npairs ? 2 * ((npairs - cluster_id + n_clusters - 1) / n_clusters) : 0)
                         : (cluster_id < npairs * a.tiles_n
                                ? (npairs * a.tiles_n - cluster_id + n_clusters - 1) / n_clusters
                                : 0);
  auto tile_of = [&](int j, int& pr, int& nt) {
    if (kLN) {
      pr = cluster_id + (j >> 1) * n_clusters;
      nt = j & 1;
    } else {
      const int tile = cluster_id + j * n_clusters;
      pr = tile / a.tiles_n;
      nt = tile - pr * a.tiles_n;
    }
  };

  if (warp == kEpiWarps) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int j = 0; j < n_mine; ++j) {
        int pr, nt;
        tile_of(j, pr, nt);
        const int mt = 2 * pr + static_cast<int>(rank);
        for (int kb = 0; kb < a.num_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&full[stage], 2u * kStageBytes);
          else mbar_arrive_remote(&full[stage], 0);
          uint8_t* st = smem + stage * kStageBytes;
          tma_load_2d_pair(st, &tmap_a, &full[stage], kb * kEncBK, mt * kEncBM);
          tma_load_2d_pair(st + kABytes, &tmap_w, &full[stage], kb * kEncBK,
                           nt * kEncBN + static_cast<int>(rank) * kWHalfRows);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == kEpiWarps + 1) {
    // ================= MMA issuer: this warp of the leader CTA =================
    // All 32 lanes run the loop convergently and one elected lane issues: with warp-uniform control
    // flow the descriptors live in uniform registers and each tcgen05.mma is a single predicated
    // instruction.  (From a lone lane inside a divergent branch the compiler wraps every MMA in an
    // elect / broadcast loop and the issuing thread, with only 384 tensor cycles per k-step to hide
    // behind, becomes the limiter.)
    if (leader) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * kEncBM, kEncBN);     // M = 256 across the pair
      uint32_t stage = 0, phase = 0, it = 0;
      const uint32_t s_addr = smem_u32(smem);
      const int num_kb = a.num_kb;
      for (int j = 0; j < n_mine; ++j, ++it) {
        const uint32_t acc = it & 1;
        mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccStride;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          if (elect_one()) {
            const uint64_t adesc = umma_desc_sw128(s_addr + stage * kStageBytes);
            const uint64_t bdesc = umma_desc_sw128(s_addr + stage * kStageBytes + kABytes);
#pragma unroll
            for (int k = 0; k < kEncBK / 16; ++k)
              umma_bf16_ss_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            umma_commit_pair(&empty[stage]);
            if (kb == num_kb - 1) umma_commit_pair(&t_full[acc]);
          }
          __syncwarp();
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ================= epilogue: bias (+ GELU | + residual [+ LayerNorm]) -> bf16 -> TMA store =================
    const int quarter = warp & 3;            // TMEM lanes [32*quarter, +32): the only ones this warp may read
    const int cg = warp >> 2;                // columns [64*cg, +64) of the tile
    uint8_t* box = boxes + warp * kBoxesPerWarp * kBoxBytes;
    uint8_t* my_row = box + lane * 128;      // this thread's token: 64 bf16, 16-byte chunks XOR-swizzled by row
    const int sw = lane & 7;
    uint32_t it = 0;
    if constexpr (kLN) {
      // One unit = 256 tokens x 384 features = both accumulators.  Pass 1 (per feature tile): accumulator +
      // bias + residual -> fp32 row sums, bf16 value left in the tile's box.  The three warps that share a
      // lane quarter exchange their partial sums through shared memory (named barrier 1 + quarter), then
      // pass 2 normalises the boxes in place and two TMA stores write the rows out — over the residual rows
      // they came from: the LayerNorm output IS the next residual stream.
      for (int j = 0; j < n_mine; j += 2) {
        int pr, nt_unused;
        tile_of(j, pr, nt_unused);
        const uint32_t unit = static_cast<uint32_t>(j) >> 1;
        const int row0 = (2 * pr + static_cast<int>(rank)) * kEncBM + quarter * 32;
        if (lane == 0) {
          bulk_wait_read_all();              // the previous unit's stores no longer read the boxes
          mbar_arrive_expect_tx(&res_bar[warp], 2 * kBoxBytes);
          tma_load_2d(box, &tmap_res, &res_bar[warp], cg * 64, row0);
          tma_load_2d(box + kBoxBytes, &tmap_res, &res_bar[warp], kEncBN + cg * 64, row0);
        }
        __syncwarp();
        float2 s1p = make_float2(0.f, 0.f), s2p = s1p;      // even / odd column partial sums
#pragma unroll 1
        for (int nt = 0; nt < 2; ++nt, ++it) {
          const uint32_t acc = it & 1;
          const int col0 = nt * kEncBN + cg * 64;
          uint8_t* row = my_row + nt * kBoxBytes;
          mbar_wait(&t_full[acc], (it >> 1) & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * kAccStride + cg * 64;
          float acc_v[2][32];
          tmem_ld_32x32(taddr, acc_v[0]);
          tmem_ld_32x32(taddr + 32, acc_v[1]);
          tmem_ld_wait();
          tc_fence_before();
          mbar_arrive_remote(&t_empty[acc], 0);
          if (nt == 0) mbar_wait(&res_bar[warp], unit & 1);
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            float(&v)[32] = acc_v[hlf];
            const float4* b4 = reinterpret_cast<const float4*>(a.bias + col0 + hlf * 32);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint4* slot = reinterpret_cast<uint4*>(row + (((hlf * 4 + c) ^ sw) << 4));
              const uint4 r = *slot;
              const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
              const float4 ba = __ldg(b4 + 2 * c), bb = __ldg(b4 + 2 * c + 1);
              const float bv[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
              uint32_t w[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[e]));
                const float2 y = fadd2(fadd2(make_float2(v[8 * c + 2 * e], v[8 * c + 2 * e + 1]),
                                             make_float2(bv[2 * e], bv[2 * e + 1])), f);
                s1p = fadd2(s1p, y);
                s2p = ffma2(y, y, s2p);
                const __nv_bfloat162 h = __floats2bfloat162_rn(y.x, y.y);
                w[e] = *reinterpret_cast<const uint32_t*>(&h);
              }
              *slot = make_uint4(w[0], w[1], w[2], w[3]);
            }
          }
        }
        // row statistics over the 384 features: this warp saw 128 of them
        float2* st = ln_stats + (unit & 1) * (kEpiWarps * 32);
        st[warp * 32 + lane] = make_float2(s1p.x + s1p.y, s2p.x + s2p.y);
        asm volatile("bar.sync %0, 96;" ::"r"(1 + quarter) : "memory");
        const float2 p0 = st[quarter * 32 + lane], p1 = st[(quarter + 4) * 32 + lane], p2 = st[(quarter + 8) * 32 + lane];
        const float inv_n = 1.0f / static_cast<float>(2 * kEncBN);
        const float mu = (p0.x + p1.x + p2.x) * inv_n;
        const float var = fmaxf((p0.y + p1.y + p2.y) * inv_n - mu * mu, 0.f);
        const float rstd = rsqrtf(var + a.ln_eps);
#pragma unroll 1
        for (int nt = 0; nt < 2; ++nt) {
          const int col0 = nt * kEncBN + cg * 64;
          uint8_t* row = my_row + nt * kBoxBytes;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint4* slot = reinterpret_cast<uint4*>(row + ((c ^ sw) << 4));
            const uint4 r = *slot;
            const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
            const float4 ga = __ldg(reinterpret_cast<const float4*>(a.ln_g + col0 + 8 * c)),
                         gb = __ldg(reinterpret_cast<const float4*>(a.ln_g + col0 + 8 * c + 4));
            const float4 ba = __ldg(reinterpret_cast<const float4*>(a.ln_b + col0 + 8 * c)),
                         bb = __ldg(reinterpret_cast<const float4*>(a.ln_b + col0 + 8 * c + 4));
            const float gv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
            const float bv[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
            uint32_t w[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              // (f - mu) rstd g + b  =  f (rstd g) + (b - mu rstd g)
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[e]));
              const float2 rg = fmul2(make_float2(gv[2 * e], gv[2 * e + 1]), splat(rstd));
              const float2 o = ffma2(f, rg, ffma2(rg, splat(-mu), make_float2(bv[2 * e], bv[2 * e + 1])));
              const __nv_bfloat162 h = __floats2bfloat162_rn(o.x, o.y);
              w[e] = *reinterpret_cast<const uint32_t*>(&h);
            }
            *slot = make_uint4(w[0], w[1], w[2], w[3]);
          }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmap_out, box, cg * 64, row0);
          tma_store_2d(&tmap_out, box + kBoxBytes, kEncBN + cg * 64, row0);
          bulk_commit_group();
        }
      }
    } else {
    for (int j = 0; j < n_mine; ++j, ++it) {
      int pr, nt;
      tile_of(j, pr, nt);
      const int mt = 2 * pr + static_cast<int>(rank);
      const int row0 = mt * kEncBM + quarter * 32;
      const int col0 = nt * kEncBN + cg * 64;
      const uint32_t acc = it & 1;
      if (kEpi == kEpiResidual) {
        if (lane == 0) {
          bulk_wait_read_all();              // the previous tile's store no longer reads the box
          mbar_arrive_expect_tx(&res_bar[warp], kBoxBytes);
          tma_load_2d(box, &tmap_res, &res_bar[warp], col0, row0);
        }
        __syncwarp();
      }
      mbar_wait(&t_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * kAccStride + cg * 64;
      bool res_ready = kEpi != kEpiResidual;
      // Bias / residual epilogues drain all 64 accumulator columns into registers and hand the accumulator
      // back at once (the MMAs of the tile after next wait for all 24 epilogue warps of the pair).  The GELU
      // epilogue is bound by instruction issue instead: it keeps only 32 columns live so that the compiler
      // can interleave more polynomial chains, and releases after the second load.
      constexpr bool kDrainFirst = kEpi != kEpiGelu;
      float acc_v[2][32];
      if (kDrainFirst) {
        tmem_ld_32x32(taddr, acc_v[0]);
        tmem_ld_32x32(taddr + 32, acc_v[1]);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive_remote(&t_empty[acc], 0);    // the leader's MMA thread owns the accumulator hand-off
      }
#pragma unroll
      for (int hlf = 0; hlf < 2; ++hlf) {
        float(&v)[32] = acc_v[hlf];
        if (!kDrainFirst) {
          tmem_ld_32x32(taddr + hlf * 32, v);
          tmem_ld_wait();
          if (hlf == 1) {
            tc_fence_before();
            mbar_arrive_remote(&t_empty[acc], 0);
          }
        }
        const float4* b4 = reinterpret_cast<const float4*>(a.bias + col0 + hlf * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b = __ldg(b4 + j);
          float2 lo = fadd2(make_float2(v[4 * j + 0], v[4 * j + 1]), make_float2(b.x, b.y));
          float2 hi = fadd2(make_float2(v[4 * j + 2], v[4 * j + 3]), make_float2(b.z, b.w));
          if (kEpi == kEpiGelu) {
            lo = gelu_erf2(lo);
            hi = gelu_erf2(hi);
          }
          v[4 * j + 0] = lo.x;
          v[4 * j + 1] = lo.y;
          v[4 * j + 2] = hi.x;
          v[4 * j + 3] = hi.y;
        }
        if (!res_ready) {
          mbar_wait(&res_bar[warp], it & 1);
          res_ready = true;
        }
        if (kEpi != kEpiResidual && hlf == 0) {
          // first write into the box: the previous tile's store (issued a whole tile ago) must have
          // finished reading it — in steady state this does not wait
          if (lane == 0) bulk_wait_read_all();
          __syncwarp();
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint4* slot = reinterpret_cast<uint4*>(my_row + (((hlf * 4 + c) ^ sw) << 4));
          if (kEpi == kEpiResidual) {
            const uint4 r = *slot;
            const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 f = fadd2(make_float2(v[8 * c + 2 * e], v[8 * c + 2 * e + 1]),
                                     __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e])));
              v[8 * c + 2 * e] = f.x;
              v[8 * c + 2 * e + 1] = f.y;
            }
          }
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[8 * c + 2 * e], v[8 * c + 2 * e + 1]);
            w[e] = *reinterpret_cast<const uint32_t*>(&h);
          }
          *slot = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(&tmap_out, box, col0, row0);
        bulk_commit_group();
      }
    }
    }
    if (lane == 0) bulk_wait_all();
  }

  __syncwarp();                                // the single-lane roles rejoin their warps
  tc_fence_before();
  cluster_sync_all();                          // nobody frees TMEM / exits while the peer still uses it
  if (warp == kEpiWarps + 1) tmem_dealloc2(tmem_base, kTmemCols);
}

template <int kEpi>
cudaError_t set_attr() {
  return cudaFuncSetAttribute(qre_enc_gemm_kernel<kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_total(kEpi));
}

}  // namespace

cudaError_t enc_gemm_init() {
  cudaError_t e = set_attr<kEpiBias>();
  if (e == cudaSuccess) e = set_attr<kEpiGelu>();
  if (e == cudaSuccess) e = set_attr<kEpiResidual>();
  if (e == cudaSuccess) e = set_attr<kEpiResidualLN>();
  return e;
}

int launch_enc_gemm(const CUtensorMap& tmap_a, const CUtensorMap& tmap_w, const CUtensorMap& tmap_out,
                    const CUtensorMap* tmap_res, const EncGemmArgs& a, int epilogue, int num_sms, cudaStream_t st,
                    bool after_kernel) {
  const int total = ((a.tiles_m + 1) / 2) * a.tiles_n;          // tiles of 256 tokens x 192 features
  if (total <= 0) return 0;
  const int clusters = num_sms / 2;
  const int work = epilogue == kEpiResidualLN ? (a.tiles_m + 1) / 2 : total;     // LayerNorm: whole 256-token units
  const int grid = 2 * (work < clusters ? work : clusters);
  const CUtensorMap& res = tmap_res != nullptr ? *tmap_res : tmap_out;
  const dim3 g(grid), blk(kThreads);
  switch (epilogue) {
    case kEpiGelu:
      launch_chain(qre_enc_gemm_kernel<kEpiGelu>, g, blk, smem_total(kEpiGelu), st, after_kernel, tmap_a, tmap_w, tmap_out,
                   res, a);
      break;
    case kEpiResidual:
      launch_chain(qre_enc_gemm_kernel<kEpiResidual>, g, blk, smem_total(kEpiResidual), st, after_kernel, tmap_a, tmap_w,
                   tmap_out, res, a);
      break;
    case kEpiResidualLN:      // caller guarantees tiles_n == 2 (384 features) and ln_g / ln_b set
      launch_chain(qre_enc_gemm_kernel<kEpiResidualLN>, g, blk, smem_total(kEpiResidualLN), st, after_kernel, tmap_a,
                   tmap_w, tmap_out, res, a);
      break;
    default:
      launch_chain(qre_enc_gemm_kernel<kEpiBias>, g, blk, smem_total(kEpiBias), st, after_kernel, tmap_a, tmap_w, tmap_out,
                   res, a);
      break;
  }
  return 1;
}

}  // namespace qre
