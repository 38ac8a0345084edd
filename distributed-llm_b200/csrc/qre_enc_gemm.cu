// qre_enc_gemm.cu — the linear layers of the on-GPU query embedder:
//   out[token, f] = epilogue( sum_k x[token, k] * W[f, k] + bias[f] )        bf16 x bf16 -> fp32
// for the four projections of a BERT layer (QKV, attention output, FFN up, FFN down), which the
// reference reaches through SentenceTransformer.encode (src/query_router_engine.py:181, 571).
//
//   * tokens are the M operand: 128 per tile, one TMEM lane per token;
//   * W is in torch.nn.Linear layout [out, in] = K-major, exactly what the UMMA B operand wants:
//     no transposition anywhere.  192 output features per tile (see qre_encoder.cuh);
//   * persistent CTAs, one per SM; a 4-stage TMA/mbarrier ring carries 128x64 of x and 192x64
//     of W per stage (40 KB); one elected thread issues tcgen05.mma (M=128, N=192, K=16) into one
//     of two TMEM accumulators, so the epilogue of tile i runs under the MMAs of tile i+1;
//   * twelve epilogue warps, one per (TMEM lane quarter, 64-column group).  Each owns a 4 KB
//     staging box in shared memory (32 tokens x 64 features, 128-byte swizzle): the residual rows
//     arrive in it by TMA while the MMAs run, the warp adds accumulator + bias (+ GELU | + residual)
//     in place, and one TMA store writes the box out — every byte of global traffic is a full
//     line moved by the copy engine; the accumulator is handed back to the MMA warp as soon as it
//     has been read into registers.
//
// Tiles are walked feature-tile-fastest, so the CTAs resident at any moment share a few token
// blocks (L2 hits on x) and all of W (<= 1.2 MB, L2-resident).
#include <stdio.h>

#include "qre_encoder.cuh"
#include "qre_ptx.cuh"

namespace qre {

namespace {

constexpr int kEpiWarps = 12;                          // warps 0-11 epilogue, 12 TMA, 13 MMA
constexpr int kThreads = (kEpiWarps + 2) * 32;
constexpr int kABytes = kEncBM * kEncBK * 2;           // 16 KB
constexpr int kBBytes = kEncBN * kEncBK * 2;           // 24 KB
constexpr int kStageBytes = kABytes + kBBytes;         // 40 KB
constexpr int kBoxBytes = 32 * 64 * 2;                 // 4 KB staging box per epilogue warp
constexpr int kSmemBar = 512;
constexpr int kSmemTotal = kEncStages * kStageBytes + kEpiWarps * kBoxBytes + kSmemBar + 1024;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kAccStride = 256;                   // TMEM column offset of the second accumulator

// GELU with the exact (erf) definition, x * Phi(x).  Phi(-a) = exp2(R(a)) for a = |x| in [0, 5.5],
// R a degree-7 polynomial fitted to log2(0.5 erfc(a / sqrt 2)) (near-minimax, Chebyshev nodes);
// Phi(a) = 1 - Phi(-a).  Measured against float64 erf over [-12, 12]: |error| <= 6.7e-7 absolute,
// 4.4e-6 relative — fp32 rounding level, two orders below the bf16 rounding that follows.  Costs
// one MUFU (ex2) and ~11 ALU instructions per element; erff costs two to three times that and the
// FFN-up epilogue is ALU-bound.
__device__ __forceinline__ float gelu_erf(float x) {
  const float a = fminf(fabsf(x), 5.5f);
  float r = fmaf(a, -1.921234571e-06f, 6.328849850e-05f);
  r = fmaf(a, r, -9.434208623e-04f);
  r = fmaf(a, r, 8.556271903e-03f);
  r = fmaf(a, r, -5.405228212e-02f);
  r = fmaf(a, r, -4.583817720e-01f);
  r = fmaf(a, r, -1.151278377e+00f);
  r = fmaf(a, r, -9.999938607e-01f);
  const float e = exp2f(r);                                // Phi(-|x|); -30 <= r <= -1, no range issues
  return x * (x < 0.f ? e : 1.0f - e);
}

template <int kEpi>
__global__ void __launch_bounds__(kThreads, 1)
qre_enc_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                    const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                    const EncGemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* boxes = smem + kEncStages * kStageBytes;                 // [kEpiWarps] x 4 KB, 1024-aligned
  uint64_t* bars = reinterpret_cast<uint64_t*>(boxes + kEpiWarps * kBoxBytes);
  uint64_t* full = bars;                        // [kEncStages] TMA -> MMA
  uint64_t* empty = bars + kEncStages;          // [kEncStages] MMA -> TMA
  uint64_t* t_full = bars + 2 * kEncStages;     // [2] MMA -> epilogue
  uint64_t* t_empty = t_full + 2;               // [2] epilogue -> MMA
  uint64_t* res_bar = t_empty + 2;              // [kEpiWarps] residual box landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kEpiWarps);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == kEpiWarps && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
    tma_prefetch_desc(&tmap_out);
    if (kEpi == kEpiResidual) tma_prefetch_desc(&tmap_res);
    for (int i = 0; i < kEncStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], kEpiWarps * 32);
    }
    for (int i = 0; i < kEpiWarps; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == kEpiWarps + 1) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total = a.tiles_m * a.tiles_n;

  if (warp == kEpiWarps) {
    // ================= TMA producer =================
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int mt = tile / a.tiles_n, nt = tile - mt * a.tiles_n;
        for (int kb = 0; kb < a.num_kb; ++kb) {
          mbar_wait(&empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full[stage], kStageBytes);
          uint8_t* st = smem + stage * kStageBytes;
          tma_load_2d(st, &tmap_a, &full[stage], kb * kEncBK, mt * kEncBM);
          tma_load_2d(st + kABytes, &tmap_w, &full[stage], kb * kEncBK, nt * kEncBN);
          if (++stage == kEncStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == kEpiWarps + 1) {
    // ================= MMA issuer (one thread) =================
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kEncBM, kEncBN);
      uint32_t stage = 0, phase = 0, it = 0;
      const uint32_t s_addr = smem_u32(smem);
      for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
        const uint32_t acc = it & 1;
        mbar_wait(&t_empty[acc], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccStride;
        for (int kb = 0; kb < a.num_kb; ++kb) {
          mbar_wait(&full[stage], phase);
          tc_fence_after();
          const uint64_t adesc = umma_desc_sw128(s_addr + stage * kStageBytes);
          const uint64_t bdesc = umma_desc_sw128(s_addr + stage * kStageBytes + kABytes);
#pragma unroll
          for (int k = 0; k < kEncBK / 16; ++k)
            umma_bf16_ss(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&empty[stage]);
          if (++stage == kEncStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&t_full[acc]);
      }
    }
  } else {
    // ================= epilogue: bias (+ GELU | + residual) -> bf16 -> TMA store =================
    const int quarter = warp & 3;            // TMEM lanes [32*quarter, +32): the only ones this warp may read
    const int cg = warp >> 2;                // columns [64*cg, +64) of the tile
    uint8_t* box = boxes + warp * kBoxBytes;
    uint8_t* my_row = box + lane * 128;      // this thread's token: 64 bf16, 16-byte chunks XOR-swizzled by row
    const int sw = lane & 7;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
      const int mt = tile / a.tiles_n, nt = tile - mt * a.tiles_n;
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
#pragma unroll 1
      for (int hlf = 0; hlf < 2; ++hlf) {
        float v[32];
        tmem_ld_32x32(taddr + hlf * 32, v);
        tmem_ld_wait();
        if (hlf == 1) {                      // accumulator fully read: the MMA warp may overwrite it
          tc_fence_before();
          mbar_arrive(&t_empty[acc]);
        }
        const float4* b4 = reinterpret_cast<const float4*>(a.bias + col0 + hlf * 32);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 b = __ldg(b4 + j);
          v[4 * j + 0] += b.x;
          v[4 * j + 1] += b.y;
          v[4 * j + 2] += b.z;
          v[4 * j + 3] += b.w;
        }
        if (kEpi == kEpiGelu) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
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
              const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
              v[8 * c + 2 * e] += f.x;
              v[8 * c + 2 * e + 1] += f.y;
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
    if (lane == 0) bulk_wait_all();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kEpiWarps + 1) tmem_dealloc(tmem_base, kTmemCols);
}

template <int kEpi>
cudaError_t set_attr() {
  return cudaFuncSetAttribute(qre_enc_gemm_kernel<kEpi>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal);
}

}  // namespace

cudaError_t enc_gemm_init() {
  cudaError_t e = set_attr<kEpiBias>();
  if (e == cudaSuccess) e = set_attr<kEpiGelu>();
  if (e == cudaSuccess) e = set_attr<kEpiResidual>();
  return e;
}

int launch_enc_gemm(const CUtensorMap& tmap_a, const CUtensorMap& tmap_w, const CUtensorMap& tmap_out,
                    const CUtensorMap* tmap_res, const EncGemmArgs& a, int epilogue, int num_sms, cudaStream_t st) {
  const int total = a.tiles_m * a.tiles_n;
  if (total <= 0) return 0;
  const int grid = total < num_sms ? total : num_sms;
  const CUtensorMap& res = tmap_res != nullptr ? *tmap_res : tmap_out;
  switch (epilogue) {
    case kEpiGelu:
      qre_enc_gemm_kernel<kEpiGelu><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, tmap_out, res, a);
      break;
    case kEpiResidual:
      qre_enc_gemm_kernel<kEpiResidual><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, tmap_out, res, a);
      break;
    default:
      qre_enc_gemm_kernel<kEpiBias><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, tmap_out, res, a);
      break;
  }
  return 1;
}

}  // namespace qre
