// qre_encoder.cu — the on-GPU query embedder behind include/qre_encoder_b200.h:
// what SentenceTransformer("all-MiniLM-L6-v2").encode does between "token ids" and "fp32
// embedding" (src/query_router_engine.py:181, 571; src/cache.py:412), restated in
// oracle/minilm_oracle.py.  This file holds the non-GEMM kernels (embedding gather + LayerNorm,
// attention, LayerNorm, mean pool + normalise), the weight/activation layout and the launch
// sequence; the linear layers are in qre_enc_gemm.cu.
//
// Activations for one batch (T packed tokens, hidden H, FFN F), all bf16, rows padded to 256:
//   x   [T, H]   layer input / residual stream          (A operand of QKV and FFN-up)
//   qkv [T, 3H]  Q | K | V column blocks                (one fused projection)
//   ctx [T, H]   attention output                       (A operand of the output projection)
//   y   [T, H]   projection + residual, before LayerNorm (unused when LayerNorm runs in the
//                projection's epilogue and the result goes straight back into x: hidden 384)
//   h   [T, F]   GELU(FFN-up)                           (A operand of FFN-down)
// LayerNorm statistics, softmax and pooling are fp32.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <new>
#include <vector>

#include "qre_encoder.cuh"
#include "qre_ptx.cuh"

using namespace qre;

namespace {

constexpr unsigned kFull = 0xFFFFFFFFu;
constexpr int kMaxHidden = 1024;
constexpr int kMaxInter = 4096;
constexpr int kMaxSeq = 512;

#define ENC_CUDA(expr)                                                                              \
  do {                                                                                              \
    cudaError_t e__ = (expr);                                                                       \
    if (e__ != cudaSuccess)                                                                         \
      return set_error(QRE_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// One warp normalises one token row held as pairs: lane owns elements 64*i + 2*lane + {0,1}.
template <int kMaxPairs>
__device__ __forceinline__ void warp_layernorm_store(float (&v)[kMaxPairs][2], int pairs, int H, const float* g,
                                                     const float* b, float eps, __nv_bfloat16* out, int lane) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPairs; ++i)
    if (i < pairs) s += v[i][0] + v[i][1];
  const float mu = warp_sum(s) / static_cast<float>(H);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPairs; ++i)
    if (i < pairs) {
      const float d0 = v[i][0] - mu, d1 = v[i][1] - mu;
      q += d0 * d0 + d1 * d1;
    }
  const float rstd = rsqrtf(warp_sum(q) / static_cast<float>(H) + eps);
#pragma unroll
  for (int i = 0; i < kMaxPairs; ++i)
    if (i < pairs) {
      const int c = 64 * i + 2 * lane;
      const float2 gg = *reinterpret_cast<const float2*>(g + c);
      const float2 bb = *reinterpret_cast<const float2*>(b + c);
      *reinterpret_cast<__nv_bfloat162*>(out + c) =
          __floats2bfloat162_rn((v[i][0] - mu) * rstd * gg.x + bb.x, (v[i][1] - mu) * rstd * gg.y + bb.y);
    }
}

constexpr int kPairsMax = kMaxHidden / 64;   // 16

// pos[t] = index of packed token t inside its sequence; one thread per sequence.
__global__ void __launch_bounds__(256)
enc_positions_kernel(const int32_t* __restrict__ offsets, int m, int max_positions, int32_t* __restrict__ pos) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int t0 = offsets[i], L = offsets[i + 1] - t0;
  // (clamped: a sequence longer than the caller promised must not index past the position table)
  for (int p = 0; p < L; ++p) pos[t0 + p] = p < max_positions ? p : max_positions - 1;
}

// x[t] = LayerNorm(word[id[t]] + position[pos[t]] + token_type[0]); one warp per packed token, so that
// the random 1.5 KB row gathers of ~100k tokens are all in flight together.  (BertEmbeddings;
// token_type_ids are all zero for one-segment input.)
__global__ void __launch_bounds__(256)
enc_embed_ln_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ tok_pos, const float* __restrict__ word,
                    const float* __restrict__ pos, const float* __restrict__ type0, const float* __restrict__ g,
                    const float* __restrict__ b, float eps, int H, int vocab, long long T, __nv_bfloat16* __restrict__ x) {
  pdl_launch_dependents();
  pdl_wait();
  const long long t = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= T) return;
  const int lane = threadIdx.x & 31;
  const int pairs = H / 64;
  int id = ids[t];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const float* wr = word + static_cast<long long>(id) * H;
  const float* pr = pos + static_cast<long long>(tok_pos[t]) * H;
  float v[kPairsMax][2];
#pragma unroll
  for (int i = 0; i < kPairsMax; ++i)
    if (i < pairs) {
      const int c = 64 * i + 2 * lane;
      const float2 a = *reinterpret_cast<const float2*>(wr + c);
      const float2 q = *reinterpret_cast<const float2*>(pr + c);
      const float2 ty = *reinterpret_cast<const float2*>(type0 + c);
      v[i][0] = a.x + q.x + ty.x;
      v[i][1] = a.y + q.y + ty.y;
    }
  warp_layernorm_store<kPairsMax>(v, pairs, H, g, b, eps, x + t * H, lane);
}

// x[t] = LayerNorm(y[t]); y already holds projection + bias + residual.  One warp per token.
__global__ void __launch_bounds__(256)
enc_layernorm_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ g, const float* __restrict__ b,
                     float eps, int H, long long T, __nv_bfloat16* __restrict__ x) {
  pdl_launch_dependents();
  pdl_wait();
  const long long t = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= T) return;
  const int lane = threadIdx.x & 31, pairs = H / 64;
  const __nv_bfloat16* yr = y + t * H;
  float v[kPairsMax][2];
#pragma unroll
  for (int i = 0; i < kPairsMax; ++i)
    if (i < pairs) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(yr + 64 * i + 2 * lane));
      v[i][0] = f.x;
      v[i][1] = f.y;
    }
  warp_layernorm_store<kPairsMax>(v, pairs, H, g, b, eps, x + t * H, lane);
}

// Same, for the common widths (H = 64 * kChunks): 8 lanes per token, four tokens per warp, 16-byte
// loads and stores, so that enough bytes are in flight per SM to run at HBM speed.
template <int kChunks>
__global__ void __launch_bounds__(256)
enc_layernorm_wide_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ g, const float* __restrict__ b,
                          float eps, long long T, __nv_bfloat16* __restrict__ x) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int H = 64 * kChunks;
  const int lane = threadIdx.x & 31, gl = lane & 7;
  const long long t = (static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4 + (lane >> 3);
  const bool ok = t < T;
  const __nv_bfloat16* yr = y + (ok ? t : 0) * H;
  float v[kChunks][8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const uint4 u = *reinterpret_cast<const uint4*>(yr + (i * 8 + gl) * 8);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[e]));
      v[i][2 * e] = f.x;
      v[i][2 * e + 1] = f.y;
      s += f.x + f.y;
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) s += __shfl_xor_sync(kFull, s, o);
  const float mu = s / static_cast<float>(H);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < kChunks; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float d = v[i][e] - mu;
      q += d * d;
    }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) q += __shfl_xor_sync(kFull, q, o);
  const float rstd = rsqrtf(q / static_cast<float>(H) + eps);
  if (!ok) return;
#pragma unroll
  for (int i = 0; i < kChunks; ++i) {
    const int c = (i * 8 + gl) * 8;
    const float4 g0 = __ldg(reinterpret_cast<const float4*>(g + c)), g1 = __ldg(reinterpret_cast<const float4*>(g + c + 4));
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(b + c)), b1 = __ldg(reinterpret_cast<const float4*>(b + c + 4));
    const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 h = __floats2bfloat162_rn((v[i][2 * e] - mu) * rstd * gg[2 * e] + bb[2 * e],
                                                     (v[i][2 * e + 1] - mu) * rstd * gg[2 * e + 1] + bb[2 * e + 1]);
      w[e] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(x + t * H + c) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

int launch_layernorm(const __nv_bfloat16* y, const float* g, const float* b, float eps, int H, long long T,
                     __nv_bfloat16* x, cudaStream_t st) {
  const dim3 wide_grid(static_cast<unsigned>((T + 31) / 32));      // 8 warps x 4 tokens per CTA
  if (H == 384)
    launch_chain(enc_layernorm_wide_kernel<6>, wide_grid, dim3(256), 0, st, true, y, g, b, eps, T, x);
  else if (H == 768)
    launch_chain(enc_layernorm_wide_kernel<12>, wide_grid, dim3(256), 0, st, true, y, g, b, eps, T, x);
  else
    launch_chain(enc_layernorm_kernel, dim3(static_cast<unsigned>((T + 7) / 8)), dim3(256), 0, st, true, y, g, b, eps, H, T,
                 x);
  return 1;
}

// ---- attention --------------------------------------------------------------------------------
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&h);
}

// ---- softmax(Q K^T / sqrt(HD)) V --------------------------------------------------------------
// The L keys and values of one (sequence, head) are staged row-major in shared memory (row stride
// HD + 8 elements: 16-byte aligned rows that fall into distinct bank groups).  A warp owns 16 query
// rows at a time and walks the keys 16 at a time with an online softmax; K and V fragments come out
// of shared memory with ldmatrix (transposed for V), scores and probabilities never leave registers
// (the accumulator layout of one m16n8k16 product is the A-operand layout of the next).
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* row_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(row_ptr))));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* row_ptr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(row_ptr))));
}

// 16-byte asynchronous copy global -> shared (LDGSTS, L2 only); src_bytes = 0 writes zeros instead.
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(
                   static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst))),
               "l"(gmem_src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// Q, K and V rows of one (sequence, head) -> shared memory with asynchronous 16-byte copies: every
// copy of the head is in flight before the first one is waited for (a register-staged loop kept
// three loads per lane in flight and the kernel ran at 2.6 TB/s, waiting on them one round trip at a
// time).  The caller waits (cp_async_wait_all) and synchronises.
template <int HD>
__device__ __forceinline__ void stage_qkv(__nv_bfloat16* Qs, __nv_bfloat16* Ks, __nv_bfloat16* Vs,
                                          const __nv_bfloat16* base, long long ld, int H, int L, int Lk, int tid,
                                          int nthreads) {
  constexpr int KS = HD + 8;
  for (int idx = tid; idx < Lk * (HD / 8); idx += nthreads) {
    const int key = idx / (HD / 8), part = idx - key * (HD / 8);
    const bool live = key < L;                                          // padded rows: zeros (keys are masked below)
    const __nv_bfloat16* p = base + (live ? key : 0) * ld + part * 8;
    const int nb = live ? 16 : 0;
    cp_async16(Qs + key * KS + part * 8, p, nb);
    cp_async16(Ks + key * KS + part * 8, p + H, nb);
    cp_async16(Vs + key * KS + part * 8, p + 2 * H, nb);
  }
}

// The context rows left in the Q slice by attend_block -> global memory, 16 bytes per thread per store.
template <int HD>
__device__ __forceinline__ void store_ctx(const __nv_bfloat16* Qs, __nv_bfloat16* out, int H, int L, int tid,
                                          int nthreads) {
  constexpr int KS = HD + 8;
  for (int idx = tid; idx < L * (HD / 8); idx += nthreads) {
    const int row = idx / (HD / 8), part = idx - row * (HD / 8);
    *reinterpret_cast<uint4*>(out + static_cast<long long>(row) * H + part * 8) =
        *reinterpret_cast<const uint4*>(Qs + row * KS + part * 8);
  }
}

// One 16-row query block of one head against the staged keys/values.  The block's context rows
// overwrite its query rows in Qs (bf16) once the query fragments are in registers.
template <int HD>
__device__ __forceinline__ void attend_block(__nv_bfloat16* Qs, const __nv_bfloat16* Ks, const __nv_bfloat16* Vs,
                                             int L, int Lk, int qb, float scale_log2e, int lane) {
  constexpr int KS = HD + 8;
  const int g = lane >> 2, t = lane & 3;
  const int mi = lane >> 3, mr = lane & 7;          // ldmatrix: this lane addresses row mr of matrix mi
  const float ninf = __int_as_float(0xff800000);
  const int r_lo = qb * 16 + g, r_hi = r_lo + 8;
  // A fragments: matrices 0..3 = rows [8 (mi & 1), +8) x dims [16 kk + 8 (mi >> 1), +8) of the block
  uint32_t qa[HD / 16][4];
#pragma unroll
  for (int kk = 0; kk < HD / 16; ++kk)
    ldmatrix_x4(qa[kk], Qs + (qb * 16 + (mi & 1) * 8 + mr) * KS + kk * 16 + (mi >> 1) * 8);
  float o[HD / 8][4];
#pragma unroll
  for (int nd = 0; nd < HD / 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f;
  float m_lo = ninf, m_hi = ninf, l_lo = 0.f, l_hi = 0.f;

  for (int k0 = 0; k0 < Lk; k0 += 16) {
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      // matrices 0..3 = keys [k0 + 8j, +8) x dims [32 k2 + 8 mi, +8): b0/b1 of two consecutive k-steps
      const __nv_bfloat16* kr = Ks + (k0 + j * 8 + mr) * KS + mi * 8;
#pragma unroll
      for (int k2 = 0; k2 < HD / 32; ++k2) {
        uint32_t kf[4];
        ldmatrix_x4(kf, kr + k2 * 32);
        mma_16816(s[j], qa[2 * k2], kf[0], kf[1]);
        mma_16816(s[j], qa[2 * k2 + 1], kf[2], kf[3]);
      }
    }
    if (k0 + 16 > L) {                    // only the last key block has padded keys to mask
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int key = k0 + j * 8 + 2 * t;
        if (key >= L) s[j][0] = s[j][2] = ninf;
        if (key + 1 >= L) s[j][1] = s[j][3] = ninf;
      }
    }
    // online softmax on the raw scores; the 1/sqrt(HD) * log2(e) factor rides in the exp2 argument
    float bm_lo = fmaxf(fmaxf(s[0][0], s[0][1]), fmaxf(s[1][0], s[1][1]));
    float bm_hi = fmaxf(fmaxf(s[0][2], s[0][3]), fmaxf(s[1][2], s[1][3]));
    bm_lo = fmaxf(bm_lo, __shfl_xor_sync(kFull, bm_lo, 1));
    bm_lo = fmaxf(bm_lo, __shfl_xor_sync(kFull, bm_lo, 2));
    bm_hi = fmaxf(bm_hi, __shfl_xor_sync(kFull, bm_hi, 1));
    bm_hi = fmaxf(bm_hi, __shfl_xor_sync(kFull, bm_hi, 2));
    const float mn_lo = fmaxf(m_lo, bm_lo), mn_hi = fmaxf(m_hi, bm_hi);   // finite: key k0 < L is in every block
    const float ms_lo = -mn_lo * scale_log2e, ms_hi = -mn_hi * scale_log2e;
    float p[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {        // exp2 arguments two at a time (FFMA2)
      const float2 a_lo = ffma2(make_float2(s[j][0], s[j][1]), splat(scale_log2e), splat(ms_lo));
      const float2 a_hi = ffma2(make_float2(s[j][2], s[j][3]), splat(scale_log2e), splat(ms_hi));
      p[j][0] = ex2_approx(a_lo.x);
      p[j][1] = ex2_approx(a_lo.y);
      p[j][2] = ex2_approx(a_hi.x);
      p[j][3] = ex2_approx(a_hi.y);
    }
    const float2 ps_a = fadd2(make_float2(p[0][0], p[0][2]), make_float2(p[0][1], p[0][3]));
    const float2 ps_b = fadd2(make_float2(p[1][0], p[1][2]), make_float2(p[1][1], p[1][3]));
    const float2 ps = fadd2(ps_a, ps_b);
    const float ps_lo = ps.x, ps_hi = ps.y;
    if (k0 > 0) {                         // rescale what earlier blocks accumulated (nothing yet in the first)
      const float c_lo = ex2_approx(fmaf(m_lo, scale_log2e, ms_lo)), c_hi = ex2_approx(fmaf(m_hi, scale_log2e, ms_hi));
      l_lo = fmaf(l_lo, c_lo, ps_lo);
      l_hi = fmaf(l_hi, c_hi, ps_hi);
#pragma unroll
      for (int nd = 0; nd < HD / 8; ++nd) {
        const float2 r_lo = fmul2(make_float2(o[nd][0], o[nd][1]), splat(c_lo));
        const float2 r_hi = fmul2(make_float2(o[nd][2], o[nd][3]), splat(c_hi));
        o[nd][0] = r_lo.x;
        o[nd][1] = r_lo.y;
        o[nd][2] = r_hi.x;
        o[nd][3] = r_hi.y;
      }
    } else {
      l_lo = ps_lo;
      l_hi = ps_hi;
    }
    m_lo = mn_lo;
    m_hi = mn_hi;
    const uint32_t pa[4] = {pack_bf16(p[0][0], p[0][1]), pack_bf16(p[0][2], p[0][3]), pack_bf16(p[1][0], p[1][1]),
                            pack_bf16(p[1][2], p[1][3])};
    // matrices 0..3 = keys [k0 + 8 (mi & 1), +8) x dims [16 n2 + 8 (mi >> 1), +8), transposed on load:
    // b0/b1 (keys 0-7 / 8-15 of the block) for two consecutive 8-dim output tiles
    const __nv_bfloat16* vr = Vs + (k0 + (mi & 1) * 8 + mr) * KS + (mi >> 1) * 8;
#pragma unroll
    for (int n2 = 0; n2 < HD / 16; ++n2) {
      uint32_t vf[4];
      ldmatrix_x4_trans(vf, vr + n2 * 16);
      mma_16816(o[2 * n2], pa, vf[0], vf[1]);
      mma_16816(o[2 * n2 + 1], pa, vf[2], vf[3]);
    }
  }
  l_lo += __shfl_xor_sync(kFull, l_lo, 1);
  l_lo += __shfl_xor_sync(kFull, l_lo, 2);
  l_hi += __shfl_xor_sync(kFull, l_hi, 1);
  l_hi += __shfl_xor_sync(kFull, l_hi, 2);
  const float i_lo = 1.f / l_lo, i_hi = 1.f / l_hi;
  __syncwarp();                                   // every lane's ldmatrix of this block's Q rows has retired
#pragma unroll
  for (int nd = 0; nd < HD / 8; ++nd) {
    const float2 n_lo = fmul2(make_float2(o[nd][0], o[nd][1]), splat(i_lo));
    const float2 n_hi = fmul2(make_float2(o[nd][2], o[nd][3]), splat(i_hi));
    *reinterpret_cast<__nv_bfloat162*>(Qs + r_lo * KS + nd * 8 + 2 * t) = __floats2bfloat162_rn(n_lo.x, n_lo.y);
    *reinterpret_cast<__nv_bfloat162*>(Qs + r_hi * KS + nd * 8 + 2 * t) = __floats2bfloat162_rn(n_hi.x, n_hi.y);
  }
}

constexpr int kShortLen = 32;     // sequences up to this long take the warp-per-head kernel

// Long sequences (kShortLen < L): one CTA per (sequence, head) at a time, its four warps share the
// staged K/V and split the query blocks.  The sequences come from `long_list` (built once per batch by
// enc_find_long_kernel; *n_long entries), walked with a grid stride so the launch stays small however
// few long sequences there are.
template <int HD>
__global__ void __launch_bounds__(128)
enc_attention_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ offsets,
                     const int32_t* __restrict__ long_list, const int32_t* __restrict__ n_long,
                     __nv_bfloat16* __restrict__ ctx, int H, int Lp, float scale_log2e) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int KS = HD + 8;
  extern __shared__ __align__(16) uint8_t att_smem[];
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(att_smem);
  __nv_bfloat16* Ks = Qs + static_cast<size_t>(Lp) * KS;
  __nv_bfloat16* Vs = Ks + static_cast<size_t>(Lp) * KS;
  const int head = blockIdx.x, n = *n_long;        // heads vary fastest in launch order: see launch_attention
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long ld = 3LL * H;
  for (int i = blockIdx.y; i < n; i += gridDim.y) {
    const int seq = long_list[i];
    const int t0 = offsets[seq];
    const int L = min(offsets[seq + 1] - t0, Lp);
    const int Lk = (L + 15) & ~15;
    const __nv_bfloat16* base = qkv + static_cast<long long>(t0) * ld + head * HD;
    __syncthreads();                               // previous sequence's rows fully copied out
    stage_qkv<HD>(Qs, Ks, Vs, base, ld, H, L, Lk, threadIdx.x, blockDim.x);
    cp_async_wait_all();
    __syncthreads();
    for (int qb = warp; qb * 16 < L; qb += blockDim.x >> 5) attend_block<HD>(Qs, Ks, Vs, L, Lk, qb, scale_log2e, lane);
    __syncthreads();
    store_ctx<HD>(Qs, ctx + static_cast<long long>(t0) * H + head * HD, H, L, threadIdx.x, blockDim.x);
  }
}

constexpr int kMidLen = 64;       // (kShortLen, kMidLen]: CTA-per-head kernel with a small shared-memory footprint

// Sorts the sequences that the warp-per-head kernel does not take into two lists: mid
// (kShortLen < L <= kMidLen) at list[0 .. n[0]) and long (L > kMidLen) at list[m .. m + n[1]).  Order within a
// list is irrelevant: every sequence is computed independently.  n[0..1] must be zero on entry.
__global__ void __launch_bounds__(256)
enc_find_long_kernel(const int32_t* __restrict__ offsets, int m, int32_t* __restrict__ list, int32_t* __restrict__ n) {
  pdl_launch_dependents();
  pdl_wait();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int L = offsets[i + 1] - offsets[i];
  if (L > kMidLen) list[m + atomicAdd(n + 1, 1)] = i;
  else if (L > kShortLen) list[atomicAdd(n, 1)] = i;
}

// Short sequences (0 < L <= kShortLen = 32, the bulk of chat queries): one WARP per (sequence, head),
// four heads of the same sequence per CTA.  Each warp stages its own head's K/V in a private slice
// of shared memory and runs the head's one or two query blocks; warps never wait for each other.  All
// global traffic is 16 bytes per thread on 64-byte row segments (Q/K/V in, context rows out).
template <int HD>
__global__ void __launch_bounds__(128)
enc_attention_short_kernel(const __nv_bfloat16* __restrict__ qkv, const int32_t* __restrict__ offsets,
                           __nv_bfloat16* __restrict__ ctx, int H, int heads, int Lp, float scale_log2e) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int KS = HD + 8;
  extern __shared__ __align__(16) uint8_t att_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t slice = static_cast<size_t>(Lp) * KS;                  // elements per Q / K / V slice
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(att_smem) + warp * 3 * slice;
  __nv_bfloat16* Ks = Qs + slice;
  __nv_bfloat16* Vs = Ks + slice;
  const int seq = blockIdx.y, head = blockIdx.x * 4 + warp;   // head groups vary fastest in launch order
  const int t0 = offsets[seq];
  const int L = offsets[seq + 1] - t0;
  if (L <= 0 || L > Lp || head >= heads) return;
  const int Lk = (L + 15) & ~15;
  const long long ld = 3LL * H;
  stage_qkv<HD>(Qs, Ks, Vs, qkv + static_cast<long long>(t0) * ld + head * HD, ld, H, L, Lk, lane, 32);
  cp_async_wait_all();
  __syncwarp();
  attend_block<HD>(Qs, Ks, Vs, L, Lk, 0, scale_log2e, lane);
  if (L > 16) attend_block<HD>(Qs, Ks, Vs, L, Lk, 1, scale_log2e, lane);
  __syncwarp();
  store_ctx<HD>(Qs, ctx + static_cast<long long>(t0) * H + head * HD, H, L, lane, 32);
}

// Pooling(mean over the sequence's tokens) + Normalize (x / max(|x|, 1e-12)); one CTA per sequence, one
// thread per pair of columns (blockDim = H / 2), four token rows in flight per thread.
__global__ void __launch_bounds__(512)
enc_pool_kernel(const __nv_bfloat16* __restrict__ x, const int32_t* __restrict__ offsets, int H, int normalize,
                float* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[16];
  const int seq = blockIdx.x;
  const int t0 = offsets[seq], L = offsets[seq + 1] - t0;
  const __nv_bfloat16* col = x + static_cast<long long>(t0) * H + 2 * threadIdx.x;
  float2 acc = make_float2(0.f, 0.f);
  int p = 0;
  for (; p + 4 <= L; p += 4) {
    float2 f[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      f[u] = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(col + static_cast<long long>(p + u) * H));
#pragma unroll
    for (int u = 0; u < 4; ++u) {            // token order, as the sequential sum of the restatement
      acc.x += f[u].x;
      acc.y += f[u].y;
    }
  }
  for (; p < L; ++p) {
    const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(col + static_cast<long long>(p) * H));
    acc.x += f.x;
    acc.y += f.y;
  }
  const float inv = 1.f / fmaxf(static_cast<float>(L), 1e-9f);
  acc.x *= inv;
  acc.y *= inv;
  float sq = warp_sum(acc.x * acc.x + acc.y * acc.y);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sq;
  __syncthreads();
  float tot = 0.f;
  for (int w = 0; w < (blockDim.x >> 5); ++w) tot += red[w];
  const float sc = normalize ? 1.f / fmaxf(sqrtf(tot), 1e-12f) : 1.f;
  *reinterpret_cast<float2*>(out + static_cast<long long>(seq) * H + 2 * threadIdx.x) = make_float2(acc.x * sc, acc.y * sc);
}

size_t attention_smem(int hd, int Lp) { return 3 * static_cast<size_t>(Lp) * (hd + 8) * 2; }   // Q, K, V slices

struct Linear {
  __nv_bfloat16* w = nullptr;   // [n, k] bf16
  float* b = nullptr;           // [n]
  CUtensorMap map;
  int n = 0, k = 0;
};

struct Layer {
  Linear qkv, out, up, down;
  float *ln1_g = nullptr, *ln1_b = nullptr, *ln2_g = nullptr, *ln2_b = nullptr;
};

inline size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

struct qre_encoder {
  qre_encoder_config_t cfg;
  int device = 0, num_sms = 0, hd = 0;
  int64_t max_tokens = 0, tok_pad = 0;
  int max_seqs = 0;
  cudaStream_t stream = nullptr;
  // attention over the longer sequences of a batch runs on `side`, beside the short-sequence kernel
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  float *word = nullptr, *pos = nullptr, *type = nullptr, *eln_g = nullptr, *eln_b = nullptr;
  std::vector<Layer> layers;
  __nv_bfloat16 *x = nullptr, *qkv = nullptr, *ctx = nullptr, *y = nullptr, *h = nullptr;
  CUtensorMap map_x, map_ctx, map_h;                       // A operands (128-token boxes)
  CUtensorMap box_qkv, box_y, box_h, box_x;                // epilogue boxes: outputs qkv / y / h, residual x
  int32_t *d_ids = nullptr, *d_off = nullptr;
  int32_t* tok_pos = nullptr;                        // [max_tokens] position of each packed token in its sequence
  int32_t *long_list = nullptr, *n_long = nullptr;   // sequences longer than kShortLen in the current batch: two buckets
  float* d_out = nullptr;
  int32_t *h_ids = nullptr, *h_off = nullptr;     // pinned staging of the host entry point
  float* h_out = nullptr;
  int64_t launches = 0;
  bool have_emb = false;
  std::vector<char> have_layer;
};

namespace {

int upload_f32(float** dst, const float* src, size_t n) {
  if (src == nullptr) return set_error(QRE_ERR_INVALID, "null weight pointer");
  if (*dst == nullptr) ENC_CUDA(cudaMalloc(dst, n * sizeof(float)));
  ENC_CUDA(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice));
  return QRE_OK;
}

// fp32 [n, k] host blocks stacked along n -> one bf16 [n_total, k] device matrix + its tensor map.
int upload_linear(Linear& L, std::initializer_list<const float*> ws, std::initializer_list<const float*> bs, int n_each,
                  int k) {
  const int parts = static_cast<int>(ws.size());
  const int n = n_each * parts;
  for (const float* p : ws)
    if (p == nullptr) return set_error(QRE_ERR_INVALID, "null weight pointer");
  for (const float* p : bs)
    if (p == nullptr) return set_error(QRE_ERR_INVALID, "null bias pointer");
  std::vector<__nv_bfloat16> wb(static_cast<size_t>(n) * k);
  std::vector<float> bb(n);
  int part = 0;
  for (const float* p : ws) {
    for (size_t i = 0; i < static_cast<size_t>(n_each) * k; ++i)
      wb[static_cast<size_t>(part) * n_each * k + i] = __float2bfloat16(p[i]);
    ++part;
  }
  part = 0;
  for (const float* p : bs) {
    memcpy(bb.data() + static_cast<size_t>(part) * n_each, p, sizeof(float) * n_each);
    ++part;
  }
  if (L.w == nullptr) {
    // rows padded to the 192-row TMA box so the last feature tile never reads past the allocation
    const size_t rows_pad = round_up(n, kEncBN);
    ENC_CUDA(cudaMalloc(&L.w, rows_pad * k * sizeof(__nv_bfloat16)));
    ENC_CUDA(cudaMemset(L.w, 0, rows_pad * k * sizeof(__nv_bfloat16)));
    ENC_CUDA(cudaMalloc(&L.b, round_up(n, kEncBN) * sizeof(float)));
    ENC_CUDA(cudaMemset(L.b, 0, round_up(n, kEncBN) * sizeof(float)));
    int rc = make_tensor_map_bf16(&L.map, L.w, static_cast<uint64_t>(k), rows_pad, kEncBK, kEncBN / 2);
    if (rc != QRE_OK) return rc;
  }
  ENC_CUDA(cudaMemcpy(L.w, wb.data(), wb.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
  ENC_CUDA(cudaMemcpy(L.b, bb.data(), bb.size() * sizeof(float), cudaMemcpyHostToDevice));
  L.n = n;
  L.k = k;
  return QRE_OK;
}

void free_linear(Linear& L) {
  cudaFree(L.w);
  cudaFree(L.b);
  L.w = nullptr;
  L.b = nullptr;
}

int run_linear(qre_encoder* e, const CUtensorMap& map_a, const Linear& L, int64_t T, const CUtensorMap& box_out,
               const CUtensorMap* box_res, int epi, cudaStream_t st, const float* ln_g = nullptr,
               const float* ln_b = nullptr, bool after_kernel = true) {
  EncGemmArgs a;
  a.num_kb = L.k / kEncBK;
  a.tiles_m = static_cast<int>((T + kEncBM - 1) / kEncBM);
  a.tiles_n = (L.n + kEncBN - 1) / kEncBN;
  a.bias = L.b;
  a.ln_g = ln_g;
  a.ln_b = ln_b;
  a.ln_eps = e->cfg.ln_eps;
  return launch_enc_gemm(map_a, L.map, box_out, box_res, a, epi, e->num_sms, st, after_kernel);
}

size_t attention_short_smem(int hd, int Lp) { return 4 * attention_smem(hd, Lp); }

// The three length buckets are independent of each other (all read the QKV projection, each writes its
// own sequences' context rows), so the two CTA-per-head launches for the longer sequences go to a
// side stream and run beside the warp-per-head kernel instead of after it: their tails (a handful of
// 100+ token sequences) no longer add to the layer.  The fork / join events cost the two neighbouring
// kernels their programmatic-dependent overlap; measured it is a clear net gain.
int launch_attention(qre_encoder* e, int m, const int32_t* offsets, int max_len, cudaStream_t st) {
  const int H = e->cfg.hidden;
  const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(e->hd));
  int n = 0;
  const __nv_bfloat16* qkv = e->qkv;
  const bool fork = max_len > kShortLen;
  if (fork) {
    // longer sequences (listed by enc_find_long_kernel): CTA per (sequence, head), once for the mid bucket
    // with a 64-row footprint (many CTAs per SM) and once, if the batch has any, for the long bucket
    ENC_CUDA(cudaEventRecord(e->ev_fork, st));
    ENC_CUDA(cudaStreamWaitEvent(e->side, e->ev_fork, 0));
    for (int bucket = 1; bucket >= 0; --bucket) {        // the long bucket first: it has the longest items
      if (max_len <= (bucket == 0 ? kShortLen : kMidLen)) continue;
      const int Lp = bucket == 0 ? (((max_len < kMidLen ? max_len : kMidLen) + 15) & ~15) : ((max_len + 15) & ~15);
      const size_t smem = attention_smem(e->hd, Lp);
      int per_sm = static_cast<int>((200u << 10) / (smem + 1024));
      per_sm = per_sm < 1 ? 1 : (per_sm > 16 ? 16 : per_sm);
      int gx = (e->num_sms * per_sm + e->cfg.heads - 1) / e->cfg.heads;
      gx = gx < 1 ? 1 : (gx > m ? m : gx);
      if (gx > 65535) gx = 65535;
      const dim3 grid(e->cfg.heads, gx);
      const int32_t* list = e->long_list + (bucket == 0 ? 0 : m);
      const int32_t* cnt = e->n_long + bucket;
      if (e->hd == 32)
        launch_chain(enc_attention_kernel<32>, grid, dim3(128), smem, e->side, false, qkv, offsets, list, cnt, e->ctx, H,
                     Lp, scale_log2e);
      else
        launch_chain(enc_attention_kernel<64>, grid, dim3(128), smem, e->side, false, qkv, offsets, list, cnt, e->ctx, H,
                     Lp, scale_log2e);
      ++n;
    }
    ENC_CUDA(cudaEventRecord(e->ev_join, e->side));
  }
  // sequences of up to kShortLen tokens: warp per head
  const int Ls = ((max_len < kShortLen ? max_len : kShortLen) + 15) & ~15;
  // x (fastest in launch order) = head group: the CTAs that read the 64-byte head slices of the same token rows
  // are resident together, so each 2304-byte qkv row is touched once while its DRAM page is open
  const dim3 grid_s((e->cfg.heads + 3) / 4, m);
  if (e->hd == 32)
    launch_chain(enc_attention_short_kernel<32>, grid_s, dim3(128), attention_short_smem(32, Ls), st, !fork, qkv,
                 offsets, e->ctx, H, e->cfg.heads, Ls, scale_log2e);
  else
    launch_chain(enc_attention_short_kernel<64>, grid_s, dim3(128), attention_short_smem(64, Ls), st, !fork, qkv,
                 offsets, e->ctx, H, e->cfg.heads, Ls, scale_log2e);
  ++n;
  if (fork) ENC_CUDA(cudaStreamWaitEvent(st, e->ev_join, 0));
  return n;
}

// The whole forward pass for m packed sequences already on the device.
int encode_on_device(qre_encoder* e, int m, const int32_t* ids, const int32_t* offsets, int64_t T, int max_len,
                     int normalize, float* out, cudaStream_t st) {
  const qre_encoder_config_t& c = e->cfg;
  const int H = c.hidden;
  bool chained = false;                  // has a kernel of this batch been launched yet (see launch_chain)
  if (T > 0) {
    if (max_len > kShortLen) {
      ENC_CUDA(cudaMemsetAsync(e->n_long, 0, 2 * sizeof(int32_t), st));
      launch_chain(enc_find_long_kernel, dim3((m + 255) / 256), dim3(256), 0, st, false, offsets, m, e->long_list,
                   e->n_long);
      chained = true;
      ++e->launches;
    }
    launch_chain(enc_positions_kernel, dim3((m + 255) / 256), dim3(256), 0, st, chained, offsets, m, c.max_positions,
                 e->tok_pos);
    launch_chain(enc_embed_ln_kernel, dim3(static_cast<unsigned>((T + 7) / 8)), dim3(256), 0, st, true, ids,
                 static_cast<const int32_t*>(e->tok_pos), static_cast<const float*>(e->word),
                 static_cast<const float*>(e->pos), static_cast<const float*>(e->type), static_cast<const float*>(e->eln_g),
                 static_cast<const float*>(e->eln_b), c.ln_eps, H, c.vocab_size, static_cast<long long>(T), e->x);
    ++e->launches;
    chained = true;
    ++e->launches;
    for (int l = 0; l < c.layers; ++l) {
      Layer& L = e->layers[l];
      e->launches += run_linear(e, e->map_x, L.qkv, T, e->box_qkv, nullptr, kEpiBias, st);
      e->launches += launch_attention(e, m, offsets, max_len, st);
      const bool joined = max_len > kShortLen;    // the stream has just waited for the side stream: plain launch next
      if (H == 2 * kEncBN) {
        // hidden 384 = two feature tiles = the two TMEM accumulators: LayerNorm runs in the projection's
        // epilogue and the result replaces the residual rows of x it was computed from
        e->launches += run_linear(e, e->map_ctx, L.out, T, e->box_x, &e->box_x, kEpiResidualLN, st, L.ln1_g, L.ln1_b,
                                  !joined);
        e->launches += run_linear(e, e->map_x, L.up, T, e->box_h, nullptr, kEpiGelu, st);
        if (2 * ((T + 255) / 256) >= e->num_sms / 2) {
          e->launches += run_linear(e, e->map_h, L.down, T, e->box_x, &e->box_x, kEpiResidualLN, st, L.ln2_g, L.ln2_b);
        } else {
          // few tokens (online batches): the fused form would walk both feature tiles' 24 k-steps each on one
          // CTA pair, one after the other; separate tiles run side by side on two pairs, LayerNorm after
          e->launches += run_linear(e, e->map_h, L.down, T, e->box_y, &e->box_x, kEpiResidual, st);
          e->launches += launch_layernorm(e->y, L.ln2_g, L.ln2_b, c.ln_eps, H, T, e->x, st);
        }
      } else {
        e->launches += run_linear(e, e->map_ctx, L.out, T, e->box_y, &e->box_x, kEpiResidual, st, nullptr, nullptr,
                                  !joined);
        e->launches += launch_layernorm(e->y, L.ln1_g, L.ln1_b, c.ln_eps, H, T, e->x, st);
        e->launches += run_linear(e, e->map_x, L.up, T, e->box_h, nullptr, kEpiGelu, st);
        e->launches += run_linear(e, e->map_h, L.down, T, e->box_y, &e->box_x, kEpiResidual, st);
        e->launches += launch_layernorm(e->y, L.ln2_g, L.ln2_b, c.ln_eps, H, T, e->x, st);
      }
    }
  }
  launch_chain(enc_pool_kernel, dim3(m), dim3(H / 2), 0, st, chained, static_cast<const __nv_bfloat16*>(e->x), offsets, H,
               normalize, out);
  ++e->launches;
  ENC_CUDA(cudaGetLastError());
  return QRE_OK;
}

int check_ready(const qre_encoder* e) {
  if (e == nullptr) return set_error(QRE_ERR_INVALID, "null encoder handle");
  if (!e->have_emb) return set_error(QRE_ERR_INVALID, "encoder: embeddings not set (qre_encoder_set_embeddings)");
  for (size_t l = 0; l < e->have_layer.size(); ++l)
    if (!e->have_layer[l]) return set_error(QRE_ERR_INVALID, "encoder: layer %d not set (qre_encoder_set_layer)", (int)l);
  return QRE_OK;
}

void destroy(qre_encoder* e) {
  if (e == nullptr) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (float* p : {e->word, e->pos, e->type, e->eln_g, e->eln_b, e->d_out}) cudaFree(p);
  for (Layer& L : e->layers) {
    free_linear(L.qkv);
    free_linear(L.out);
    free_linear(L.up);
    free_linear(L.down);
    for (float* p : {L.ln1_g, L.ln1_b, L.ln2_g, L.ln2_b}) cudaFree(p);
  }
  for (__nv_bfloat16* p : {e->x, e->qkv, e->ctx, e->y, e->h}) cudaFree(p);
  cudaFree(e->d_ids);
  cudaFree(e->d_off);
  cudaFree(e->tok_pos);
  cudaFree(e->long_list);
  cudaFree(e->n_long);
  cudaFreeHost(e->h_ids);
  cudaFreeHost(e->h_off);
  cudaFreeHost(e->h_out);
  if (e->stream) cudaStreamDestroy(e->stream);
  if (e->side) cudaStreamDestroy(e->side);
  if (e->ev_fork) cudaEventDestroy(e->ev_fork);
  if (e->ev_join) cudaEventDestroy(e->ev_join);
  delete e;
}

}  // namespace

extern "C" {

int qre_encoder_create(int device, const qre_encoder_config_t* cfg, int64_t max_tokens, int max_seqs,
                       qre_encoder_t** out) {
  if (cfg == nullptr || out == nullptr) return set_error(QRE_ERR_INVALID, "null argument");
  *out = nullptr;
  const qre_encoder_config_t& c = *cfg;
  if (c.hidden <= 0 || c.hidden % 64 != 0 || c.hidden > kMaxHidden)
    return set_error(QRE_ERR_INVALID, "encoder: hidden must be a multiple of 64 and <= %d (got %d)", kMaxHidden, c.hidden);
  if (c.intermediate <= 0 || c.intermediate % 64 != 0 || c.intermediate > kMaxInter)
    return set_error(QRE_ERR_INVALID, "encoder: intermediate must be a multiple of 64 and <= %d (got %d)", kMaxInter,
                     c.intermediate);
  if (c.heads <= 0 || c.hidden % c.heads != 0 || (c.hidden / c.heads != 32 && c.hidden / c.heads != 64))
    return set_error(QRE_ERR_INVALID, "encoder: hidden / heads must be 32 or 64 (got %d / %d)", c.hidden, c.heads);
  if (c.layers <= 0 || c.layers > 48 || c.vocab_size <= 0 || c.type_vocab <= 0)
    return set_error(QRE_ERR_INVALID, "encoder: bad layers / vocab_size / type_vocab");
  if (c.max_seq_len <= 0 || c.max_seq_len > kMaxSeq || c.max_seq_len > c.max_positions)
    return set_error(QRE_ERR_INVALID, "encoder: max_seq_len must be in [1, min(%d, max_positions)]", kMaxSeq);
  if (!(c.ln_eps >= 0.f)) return set_error(QRE_ERR_INVALID, "encoder: bad ln_eps");
  if (max_tokens <= 0 || max_seqs <= 0 || max_tokens > (1LL << 30) || max_seqs > 65535)
    return set_error(QRE_ERR_INVALID, "encoder: need 0 < max_tokens <= 2^30 and 0 < max_seqs <= 65535");
  int n_dev = 0;
  ENC_CUDA(cudaGetDeviceCount(&n_dev));
  if (device < 0 || device >= n_dev) return set_error(QRE_ERR_NO_DEVICE, "encoder: no CUDA device %d", device);
  ENC_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  ENC_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error(QRE_ERR_NO_DEVICE, "encoder: device %d is sm_%d%d; this library is built for sm_100a only", device,
                     prop.major, prop.minor);
  qre_encoder* e = new (std::nothrow) qre_encoder();
  if (e == nullptr) return set_error(QRE_ERR_CUDA, "out of host memory");
  e->cfg = c;
  e->device = device;
  e->num_sms = prop.multiProcessorCount;
  e->hd = c.hidden / c.heads;
  e->max_tokens = max_tokens;
  e->tok_pad = static_cast<int64_t>(round_up(static_cast<size_t>(max_tokens), kEncRowPad));
  e->max_seqs = max_seqs;
  e->layers.resize(c.layers);
  e->have_layer.assign(c.layers, 0);
  auto fail_free = [&](int rc) {
    destroy(e);
    return rc;
  };
#define ENC_TRY(expr)                                                                                \
  do {                                                                                               \
    cudaError_t e__ = (expr);                                                                        \
    if (e__ != cudaSuccess)                                                                          \
      return fail_free(set_error(e__ == cudaErrorMemoryAllocation ? QRE_ERR_CUDA : QRE_ERR_CUDA,    \
                                 "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__)); \
  } while (0)
  ENC_TRY(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  ENC_TRY(cudaStreamCreateWithFlags(&e->side, cudaStreamNonBlocking));
  ENC_TRY(cudaEventCreateWithFlags(&e->ev_fork, cudaEventDisableTiming));
  ENC_TRY(cudaEventCreateWithFlags(&e->ev_join, cudaEventDisableTiming));
  ENC_TRY(enc_gemm_init());
  const size_t att_s = attention_short_smem(e->hd, kShortLen);
  if (att_s > 48 * 1024) {
    if (e->hd == 32)
      ENC_TRY(cudaFuncSetAttribute(enc_attention_short_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_s));
    else
      ENC_TRY(cudaFuncSetAttribute(enc_attention_short_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att_s));
  }
  const size_t att = attention_smem(e->hd, (c.max_seq_len + 15) & ~15);
  if (att > 48 * 1024) {
    if (e->hd == 32)
      ENC_TRY(cudaFuncSetAttribute(enc_attention_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att));
    else
      ENC_TRY(cudaFuncSetAttribute(enc_attention_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)att));
  }
  const size_t T = static_cast<size_t>(e->tok_pad), H = c.hidden, F = c.intermediate;
  ENC_TRY(cudaMalloc(&e->x, T * H * 2));
  ENC_TRY(cudaMalloc(&e->qkv, T * 3 * H * 2));
  ENC_TRY(cudaMalloc(&e->ctx, T * H * 2));
  ENC_TRY(cudaMalloc(&e->y, T * H * 2));
  ENC_TRY(cudaMalloc(&e->h, T * F * 2));
  // rows past the batch's token count are read by the last TMA box of a GEMM (never stored):
  // keep them finite
  ENC_TRY(cudaMemset(e->x, 0, T * H * 2));
  ENC_TRY(cudaMemset(e->ctx, 0, T * H * 2));
  ENC_TRY(cudaMemset(e->h, 0, T * F * 2));
  ENC_TRY(cudaMalloc(&e->d_ids, sizeof(int32_t) * max_tokens));
  ENC_TRY(cudaMalloc(&e->d_off, sizeof(int32_t) * (static_cast<size_t>(max_seqs) + 1)));
  ENC_TRY(cudaMalloc(&e->d_out, sizeof(float) * static_cast<size_t>(max_seqs) * H));
  ENC_TRY(cudaMalloc(&e->tok_pos, sizeof(int32_t) * static_cast<size_t>(max_tokens)));
  ENC_TRY(cudaMalloc(&e->long_list, 2 * sizeof(int32_t) * static_cast<size_t>(max_seqs)));
  ENC_TRY(cudaMalloc(&e->n_long, 2 * sizeof(int32_t)));
  ENC_TRY(cudaMallocHost(&e->h_ids, sizeof(int32_t) * max_tokens));
  ENC_TRY(cudaMallocHost(&e->h_off, sizeof(int32_t) * (static_cast<size_t>(max_seqs) + 1)));
  ENC_TRY(cudaMallocHost(&e->h_out, sizeof(float) * static_cast<size_t>(max_seqs) * H));
#undef ENC_TRY
  int rc = make_tensor_map_bf16(&e->map_x, e->x, H, T, kEncBK, kEncBM);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->map_ctx, e->ctx, H, T, kEncBK, kEncBM);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->map_h, e->h, F, T, kEncBK, kEncBM);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->box_qkv, e->qkv, 3 * H, T, kEncBoxCols, kEncBoxRows);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->box_y, e->y, H, T, kEncBoxCols, kEncBoxRows);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->box_h, e->h, F, T, kEncBoxCols, kEncBoxRows);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&e->box_x, e->x, H, T, kEncBoxCols, kEncBoxRows);
  if (rc != QRE_OK) return fail_free(rc);
  *out = e;
  return QRE_OK;
}

int qre_encoder_destroy(qre_encoder_t* e) {
  destroy(e);
  return QRE_OK;
}

int qre_encoder_set_embeddings(qre_encoder_t* e, const float* word, const float* position, const float* token_type,
                               const float* ln_g, const float* ln_b) {
  if (e == nullptr) return set_error(QRE_ERR_INVALID, "null encoder handle");
  ENC_CUDA(cudaSetDevice(e->device));
  ENC_CUDA(cudaStreamSynchronize(e->stream));
  const size_t H = e->cfg.hidden;
  int rc = upload_f32(&e->word, word, static_cast<size_t>(e->cfg.vocab_size) * H);
  if (rc == QRE_OK) rc = upload_f32(&e->pos, position, static_cast<size_t>(e->cfg.max_positions) * H);
  if (rc == QRE_OK) rc = upload_f32(&e->type, token_type, static_cast<size_t>(e->cfg.type_vocab) * H);
  if (rc == QRE_OK) rc = upload_f32(&e->eln_g, ln_g, H);
  if (rc == QRE_OK) rc = upload_f32(&e->eln_b, ln_b, H);
  if (rc == QRE_OK) e->have_emb = true;
  return rc;
}

int qre_encoder_set_layer(qre_encoder_t* e, int layer, const qre_encoder_layer_t* w) {
  if (e == nullptr || w == nullptr) return set_error(QRE_ERR_INVALID, "null argument");
  if (layer < 0 || layer >= e->cfg.layers) return set_error(QRE_ERR_INVALID, "encoder: layer %d out of range", layer);
  ENC_CUDA(cudaSetDevice(e->device));
  ENC_CUDA(cudaStreamSynchronize(e->stream));
  Layer& L = e->layers[layer];
  const int H = e->cfg.hidden, F = e->cfg.intermediate;
  int rc = upload_linear(L.qkv, {w->wq, w->wk, w->wv}, {w->bq, w->bk, w->bv}, H, H);
  if (rc == QRE_OK) rc = upload_linear(L.out, {w->wo}, {w->bo}, H, H);
  if (rc == QRE_OK) rc = upload_linear(L.up, {w->w1}, {w->b1}, F, H);
  if (rc == QRE_OK) rc = upload_linear(L.down, {w->w2}, {w->b2}, H, F);
  if (rc == QRE_OK) rc = upload_f32(&L.ln1_g, w->ln1_g, H);
  if (rc == QRE_OK) rc = upload_f32(&L.ln1_b, w->ln1_b, H);
  if (rc == QRE_OK) rc = upload_f32(&L.ln2_g, w->ln2_g, H);
  if (rc == QRE_OK) rc = upload_f32(&L.ln2_b, w->ln2_b, H);
  if (rc == QRE_OK) e->have_layer[layer] = 1;
  return rc;
}

int qre_encode_dev(qre_encoder_t* e, int m, const int32_t* ids, const int32_t* offsets, int64_t n_tokens, int max_len,
                   int normalize, float* out, void* stream) {
  int rc = check_ready(e);
  if (rc != QRE_OK) return rc;
  if (m == 0) return QRE_OK;
  if (m < 0 || m > e->max_seqs) return set_error(QRE_ERR_CAPACITY, "encoder: %d sequences, capacity %d", m, e->max_seqs);
  if (n_tokens < 0 || n_tokens > e->max_tokens)
    return set_error(QRE_ERR_CAPACITY, "encoder: %lld tokens, capacity %lld", (long long)n_tokens, (long long)e->max_tokens);
  if (max_len < 0 || max_len > e->cfg.max_seq_len)
    return set_error(QRE_ERR_INVALID, "encoder: max_len %d exceeds max_seq_len %d", max_len, e->cfg.max_seq_len);
  if (offsets == nullptr || out == nullptr || (n_tokens > 0 && ids == nullptr))
    return set_error(QRE_ERR_INVALID, "null argument");
  ENC_CUDA(cudaSetDevice(e->device));
  cudaStream_t st = stream != nullptr ? static_cast<cudaStream_t>(stream) : e->stream;
  return encode_on_device(e, m, ids, offsets, n_tokens, max_len, normalize, out, st);
}

int qre_encode(qre_encoder_t* e, int m, const int32_t* ids, const int32_t* offsets, int normalize, float* out) {
  int rc = check_ready(e);
  if (rc != QRE_OK) return rc;
  if (m == 0) return QRE_OK;
  if (m < 0 || m > e->max_seqs) return set_error(QRE_ERR_CAPACITY, "encoder: %d sequences, capacity %d", m, e->max_seqs);
  if (offsets == nullptr || out == nullptr) return set_error(QRE_ERR_INVALID, "null argument");
  if (offsets[0] != 0) return set_error(QRE_ERR_INVALID, "encoder: offsets[0] must be 0");
  // pack with truncation into the pinned staging buffers
  int64_t T = 0;
  int max_len = 0;
  e->h_off[0] = 0;
  for (int i = 0; i < m; ++i) {
    const int64_t len_in = static_cast<int64_t>(offsets[i + 1]) - offsets[i];
    if (len_in < 0) return set_error(QRE_ERR_INVALID, "encoder: offsets must be non-decreasing (sequence %d)", i);
    const int len = static_cast<int>(len_in < e->cfg.max_seq_len ? len_in : e->cfg.max_seq_len);
    if (T + len > e->max_tokens)
      return set_error(QRE_ERR_CAPACITY, "encoder: more than %lld tokens in the batch", (long long)e->max_tokens);
    if (len > 0) {
      if (ids == nullptr) return set_error(QRE_ERR_INVALID, "null ids");
      memcpy(e->h_ids + T, ids + offsets[i], sizeof(int32_t) * len);
    }
    T += len;
    e->h_off[i + 1] = static_cast<int32_t>(T);
    if (len > max_len) max_len = len;
  }
  ENC_CUDA(cudaSetDevice(e->device));
  cudaStream_t st = e->stream;
  if (T > 0) ENC_CUDA(cudaMemcpyAsync(e->d_ids, e->h_ids, sizeof(int32_t) * T, cudaMemcpyHostToDevice, st));
  ENC_CUDA(cudaMemcpyAsync(e->d_off, e->h_off, sizeof(int32_t) * (m + 1), cudaMemcpyHostToDevice, st));
  rc = encode_on_device(e, m, e->d_ids, e->d_off, T, max_len, normalize, e->d_out, st);
  if (rc != QRE_OK) return rc;
  const size_t bytes = sizeof(float) * static_cast<size_t>(m) * e->cfg.hidden;
  ENC_CUDA(cudaMemcpyAsync(e->h_out, e->d_out, bytes, cudaMemcpyDeviceToHost, st));
  ENC_CUDA(cudaStreamSynchronize(st));
  memcpy(out, e->h_out, bytes);
  return QRE_OK;
}

int qre_encoder_launch_count(const qre_encoder_t* e, int64_t* count) {
  if (e == nullptr || count == nullptr) return set_error(QRE_ERR_INVALID, "null argument");
  *count = e->launches;
  return QRE_OK;
}

int qre_encoder_debug_gemm(int device, int rows, int n, int k, const float* a, const float* w, const float* bias,
                           const float* residual, int epilogue, const float* ln_g, const float* ln_b, float ln_eps,
                           float* c) {
  if (rows <= 0 || n <= 0 || k <= 0 || n % 8 != 0 || k % 64 != 0)
    return set_error(QRE_ERR_INVALID, "debug_gemm: need rows > 0, n %% 8 == 0, k %% 64 == 0");
  if (a == nullptr || w == nullptr || bias == nullptr || c == nullptr || (epilogue >= kEpiResidual && residual == nullptr))
    return set_error(QRE_ERR_INVALID, "null argument");
  if (epilogue < 0 || epilogue > 3) return set_error(QRE_ERR_INVALID, "debug_gemm: epilogue must be 0..3");
  if (epilogue == kEpiResidualLN && (n != 2 * kEncBN || ln_g == nullptr || ln_b == nullptr))
    return set_error(QRE_ERR_INVALID, "debug_gemm: the LayerNorm epilogue needs n == %d and ln_g / ln_b", 2 * kEncBN);
  ENC_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  ENC_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return set_error(QRE_ERR_NO_DEVICE, "debug_gemm: sm_100a only");
  ENC_CUDA(enc_gemm_init());
  const size_t rows_pad = round_up(rows, kEncRowPad), n_pad = round_up(n, kEncBN);
  std::vector<__nv_bfloat16> ab(rows_pad * k, __float2bfloat16(0.f)), wb(n_pad * k, __float2bfloat16(0.f)),
      rb(rows_pad * n, __float2bfloat16(0.f)), cb(rows_pad * n);
  for (size_t i = 0; i < static_cast<size_t>(rows) * k; ++i) ab[i] = __float2bfloat16(a[i]);
  for (size_t i = 0; i < static_cast<size_t>(n) * k; ++i) wb[i] = __float2bfloat16(w[i]);
  if (residual != nullptr)
    for (size_t i = 0; i < static_cast<size_t>(rows) * n; ++i) rb[i] = __float2bfloat16(residual[i]);
  __nv_bfloat16 *da = nullptr, *dw = nullptr, *dr = nullptr, *dc = nullptr;
  float *db = nullptr, *dg = nullptr, *dbeta = nullptr;
  int rc = QRE_OK;
  auto cleanup = [&]() {
    cudaFree(da);
    cudaFree(dw);
    cudaFree(dr);
    cudaFree(dc);
    cudaFree(db);
    cudaFree(dg);
    cudaFree(dbeta);
  };
#define DBG_TRY(expr)                                                                              \
  do {                                                                                             \
    cudaError_t e__ = (expr);                                                                      \
    if (e__ != cudaSuccess) {                                                                      \
      cleanup();                                                                                   \
      return set_error(QRE_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
    }                                                                                              \
  } while (0)
  DBG_TRY(cudaMalloc(&da, ab.size() * 2));
  DBG_TRY(cudaMalloc(&dw, wb.size() * 2));
  DBG_TRY(cudaMalloc(&dr, rb.size() * 2));
  DBG_TRY(cudaMalloc(&dc, cb.size() * 2));
  DBG_TRY(cudaMalloc(&db, n_pad * sizeof(float)));
  DBG_TRY(cudaMemset(db, 0, n_pad * sizeof(float)));
  DBG_TRY(cudaMemcpy(da, ab.data(), ab.size() * 2, cudaMemcpyHostToDevice));
  DBG_TRY(cudaMemcpy(dw, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
  DBG_TRY(cudaMemcpy(dr, rb.data(), rb.size() * 2, cudaMemcpyHostToDevice));
  DBG_TRY(cudaMemcpy(db, bias, n * sizeof(float), cudaMemcpyHostToDevice));
  if (epilogue == kEpiResidualLN) {
    DBG_TRY(cudaMalloc(&dg, n * sizeof(float)));
    DBG_TRY(cudaMalloc(&dbeta, n * sizeof(float)));
    DBG_TRY(cudaMemcpy(dg, ln_g, n * sizeof(float), cudaMemcpyHostToDevice));
    DBG_TRY(cudaMemcpy(dbeta, ln_b, n * sizeof(float), cudaMemcpyHostToDevice));
  }
  CUtensorMap ma, mw, mo, mr;
  rc = make_tensor_map_bf16(&ma, da, k, rows_pad, kEncBK, kEncBM);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&mw, dw, k, n_pad, kEncBK, kEncBN / 2);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&mo, dc, n, rows_pad, kEncBoxCols, kEncBoxRows);
  if (rc == QRE_OK) rc = make_tensor_map_bf16(&mr, dr, n, rows_pad, kEncBoxCols, kEncBoxRows);
  if (rc != QRE_OK) {
    cleanup();
    return rc;
  }
  EncGemmArgs g;
  g.num_kb = k / kEncBK;
  g.tiles_m = static_cast<int>(rows_pad / kEncBM);
  g.tiles_n = static_cast<int>(n_pad / kEncBN);
  g.bias = db;
  g.ln_g = dg;
  g.ln_b = dbeta;
  g.ln_eps = ln_eps;
  launch_enc_gemm(ma, mw, mo, &mr, g, epilogue, prop.multiProcessorCount, nullptr);
  DBG_TRY(cudaGetLastError());
  DBG_TRY(cudaDeviceSynchronize());
  DBG_TRY(cudaMemcpy(cb.data(), dc, cb.size() * 2, cudaMemcpyDeviceToHost));
#undef DBG_TRY
  cleanup();
  for (size_t i = 0; i < static_cast<size_t>(rows) * n; ++i) c[i] = __bfloat162float(cb[i]);
  return QRE_OK;
}

}  // extern "C"
