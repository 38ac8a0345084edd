// qre_enc_gemm.cu — the linear layers of the on-GPU query embedder:
//   out[token, f] = epilogue( sum_k x[token, k] * W[f, k] + bias[f] )        bf16 x bf16 -> fp32
// for the four projections of a BERT layer (QKV, attention output, FFN up, FFN down), which the
// reference reaches through SentenceTransformer.encode (src/query_router_engine.py:181, 571).
//
//   * tokens are the M operand: 128 per tile, one TMEM lane per token, so each epilogue thread
//     owns one token's slice of the output row and writes 64 contiguous bytes at a time;
//   * W is in torch.nn.Linear layout [out, in] = K-major, exactly what the UMMA B operand wants:
//     no transposition anywhere.  192 output features per tile (see qre_encoder.cuh);
//   * persistent CTAs, one per SM; a 5-stage TMA/mbarrier ring carries 128x64 of x and 192x64
//     of W per stage (40 KB); one elected thread issues tcgen05.mma (M=128, N=192, K=16) into one
//     of two TMEM accumulators, so the epilogue of tile i runs under the MMAs of tile i+1;
//   * eight epilogue warps (two per TMEM lane quarter, 96 columns each) add the bias and apply
//     the fused tail: nothing / exact-erf GELU / + residual row, then round to bf16.
//
// Tiles are walked feature-tile-fastest, so the CTAs resident at any moment share a few token
// blocks (L2 hits on x) and all of W (<= 1.2 MB, L2-resident).
#include <stdio.h>

#include "qre_encoder.cuh"
#include "qre_ptx.cuh"

namespace qre {

namespace {

constexpr int kThreads = 320;                          // warps 0-7 epilogue, 8 TMA, 9 MMA
constexpr int kABytes = kEncBM * kEncBK * 2;           // 16 KB
constexpr int kBBytes = kEncBN * kEncBK * 2;           // 24 KB
constexpr int kStageBytes = kABytes + kBBytes;         // 40 KB
constexpr int kSmemBar = 256;
constexpr int kSmemTotal = kEncStages * kStageBytes + kSmemBar + 1024;
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kAccStride = 256;                   // TMEM column offset of the second accumulator
constexpr int kEpiWarps = 8;
constexpr int kColsPerWarp = kEncBN / 2;               // 96

__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}

template <int kEpi>
__global__ void __launch_bounds__(kThreads, 1)
qre_enc_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                    const EncGemmArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kEncStages * kStageBytes);
  uint64_t* full = bars;                        // [kEncStages] TMA -> MMA
  uint64_t* empty = bars + kEncStages;          // [kEncStages] MMA -> TMA
  uint64_t* t_full = bars + 2 * kEncStages;     // [2] MMA -> epilogue
  uint64_t* t_empty = t_full + 2;               // [2] epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_w);
    for (int i = 0; i < kEncStages; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&t_full[i], 1);
      mbar_init(&t_empty[i], kEpiWarps * 32);
    }
    fence_barrier_init();
  }
  if (warp == 9) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total = a.tiles_m * a.tiles_n;

  if (warp == 8) {
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
  } else if (warp == 9) {
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
    // ================= epilogue: bias (+ GELU | + residual) -> bf16 =================
    const int quarter = warp & 3;            // TMEM lanes [32*quarter, +32): the only ones this warp may read
    const int half = warp >> 2;              // columns [96*half, +96) of the tile
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x, ++it) {
      const int mt = tile / a.tiles_n, nt = tile - mt * a.tiles_n;
      const uint32_t acc = it & 1;
      mbar_wait(&t_full[acc], (it >> 1) & 1);
      tc_fence_after();
      const long long row = static_cast<long long>(mt) * kEncBM + quarter * 32 + lane;
      const bool row_ok = row < a.rows;
#pragma unroll 1
      for (int c = 0; c < kColsPerWarp / 32; ++c) {
        const int col_t = half * kColsPerWarp + c * 32;
        const int col0 = nt * kEncBN + col_t;
        float v[32];
        tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) + acc * kAccStride + col_t, v);
        tmem_ld_wait();
        if (col0 < a.n) {                    // n % 32 == 0: a 32-column chunk is all in or all out
          const float4* b4 = reinterpret_cast<const float4*>(a.bias + col0);
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
          if (row_ok) {
            const long long off = row * a.n + col0;
            if (kEpi == kEpiResidual) {
              const uint4* r4 = reinterpret_cast<const uint4*>(a.residual + off);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint4 r = __ldg(r4 + j);
                const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
                  v[8 * j + 2 * e] += f.x;
                  v[8 * j + 2 * e + 1] += f.y;
                }
              }
            }
            uint4* o4 = reinterpret_cast<uint4*>(a.out + off);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint32_t w[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 h = __floats2bfloat162_rn(v[8 * j + 2 * e], v[8 * j + 2 * e + 1]);
                w[e] = *reinterpret_cast<const uint32_t*>(&h);
              }
              o4[j] = make_uint4(w[0], w[1], w[2], w[3]);
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&t_empty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, kTmemCols);
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

int launch_enc_gemm(const CUtensorMap& tmap_a, const CUtensorMap& tmap_w, const EncGemmArgs& a, int epilogue,
                    int num_sms, cudaStream_t st) {
  const int total = a.tiles_m * a.tiles_n;
  if (total <= 0) return 0;
  const int grid = total < num_sms ? total : num_sms;
  switch (epilogue) {
    case kEpiGelu:
      qre_enc_gemm_kernel<kEpiGelu><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, a);
      break;
    case kEpiResidual:
      qre_enc_gemm_kernel<kEpiResidual><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, a);
      break;
    default:
      qre_enc_gemm_kernel<kEpiBias><<<grid, kThreads, kSmemTotal, st>>>(tmap_a, tmap_w, a);
      break;
  }
  return 1;
}

}  // namespace qre
