// qre_scan_epilogue.cuh — the fused top-k epilogue shared by both scan kernels
// (qre_gemm_topk.cu single-CTA, qre_gemm_topk2.cu CTA-pair).
//
// One epilogue thread owns one query (one TMEM lane).  Per 256-column accumulator tile it reads
// 4 chunks of 64 scores with tcgen05.ld, reduces each chunk to its maximum with 3-input FMNMX
// and compares it with the query's current k-th best; only chunks that can still matter take
// the candidate path (row mask + insertion).
#pragma once
#include "qre_common.cuh"
#include "qre_ptx.cuh"

namespace qre {

struct TopK {              // per-thread candidate list; lives in local memory (rare path only)
  float v[kTopK];
  uint32_t r[kTopK];
};

// Candidate path: mask (row bound, context, TTL — src/cache.py:274-276) then insert.
// Returns the new k-th best.  Deliberately not inlined: the fast path must stay tiny.
static __device__ __noinline__ float consider(TopK* tk, float v, long long row, long long n_rows,
                                              const RowMeta* __restrict__ meta, uint64_t myctx, long long now_us,
                                              long long ttl_us) {
  if (row >= n_rows) return tk->v[kTopK - 1];
  const int4 raw = __ldg(reinterpret_cast<const int4*>(meta + row));
  const uint64_t ctx = (static_cast<uint64_t>(static_cast<uint32_t>(raw.y)) << 32) | static_cast<uint32_t>(raw.x);
  const long long ts = static_cast<long long>((static_cast<uint64_t>(static_cast<uint32_t>(raw.w)) << 32) |
                                              static_cast<uint32_t>(raw.z));
  if (ctx != myctx || (now_us - ts) > ttl_us) return tk->v[kTopK - 1];
  int i = kTopK - 1;
  while (i > 0 && v > tk->v[i - 1]) {      // strict: equal scores keep the earlier row first
    tk->v[i] = tk->v[i - 1];
    tk->r[i] = tk->r[i - 1];
    --i;
  }
  tk->v[i] = v;
  tk->r[i] = static_cast<uint32_t>(row);
  return tk->v[kTopK - 1];
}

__device__ __forceinline__ void topk_reset(TopK& tk, float floor_v) {
#pragma unroll
  for (int i = 0; i < kTopK; ++i) { tk.v[i] = floor_v; tk.r[i] = kNoRow; }
}

__device__ __forceinline__ void topk_store(const TopK& tk, PartialEntry* out) {
#pragma unroll
  for (int i = 0; i < kTopK; ++i) out[i] = PartialEntry{tk.v[i], tk.r[i]};
}

// Scan one accumulator tile (this thread's lane, kBN columns starting at `taddr`); returns the
// updated k-th best.  kDump additionally writes the raw scores out (bring-up / tests).
template <bool kDump>
__device__ __forceinline__ float scan_tile(uint32_t taddr, long long row_base, int q, bool active, float kth,
                                           TopK& tk, uint64_t myctx, const GemmArgs& a) {
#pragma unroll 1
  for (int c = 0; c < kBN / 64; ++c) {
    float v[64];
    tmem_ld_32x32(taddr + c * 64, v);
    tmem_ld_32x32(taddr + c * 64 + 32, v + 32);
    tmem_ld_wait();
    if (kDump) {
      if (active && a.dump != nullptr) {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const long long row = row_base + c * 64 + i;
          if (row >= a.dump_row0 && row < a.dump_row0 + a.dump_ld && row < a.n_rows)
            a.dump[static_cast<long long>(q) * a.dump_ld + (row - a.dump_row0)] = v[i];
        }
      }
    }
    float m0 = fmaxf(v[0], v[1]), m1 = fmaxf(v[2], v[3]), m2 = fmaxf(v[4], v[5]), m3 = fmaxf(v[6], v[7]);
#pragma unroll
    for (int i = 8; i < 64; i += 8) {
      m0 = fmaxf(m0, fmaxf(v[i], v[i + 1]));
      m1 = fmaxf(m1, fmaxf(v[i + 2], v[i + 3]));
      m2 = fmaxf(m2, fmaxf(v[i + 4], v[i + 5]));
      m3 = fmaxf(m3, fmaxf(v[i + 6], v[i + 7]));
    }
    const float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
    if (mx > kth) {
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        if (v[i] > kth)
          kth = consider(&tk, v[i], row_base + c * 64 + i, a.n_rows, a.meta, myctx, a.now_us, a.ttl_us);
      }
    }
  }
  return kth;
}

}  // namespace qre
