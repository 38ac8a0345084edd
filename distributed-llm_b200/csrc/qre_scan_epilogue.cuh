// qre_scan_epilogue.cuh — the fused top-k epilogue shared by both scan kernels
// (qre_gemm_topk.cu single-CTA, qre_gemm_topk2.cu CTA-pair).
//
// One epilogue thread owns one query (one TMEM lane).  Per 256-column accumulator tile it reads
// 4 chunks of 64 scores with tcgen05.ld, reduces each chunk to its maximum with 3-input FMNMX
// and compares it with the query's current gate; only chunks that can still matter take the
// candidate path (row mask + insertion).
//
// Gate.  The exact winner of a query has approx >= a* - 2 eps where a* is the
// best approximate score over all masked rows; nothing under the hit floor can hit.  Within one
// (query, row range) the thread keeps the 8 best approximate scores in a sorted list plus a
// chain of spill chunks for candidates that do not fit but are still inside the window of the
// running best.  A score is looked at iff
//     v >= gate,   gate = min( list[7],  max(floor, list[0] - 2 eps) )
// (while the list is not full list[7] is the floor).  Nothing with v >= final window is ever
// dropped: it is in the list or in the chain, so the re-score pass is exact without a second
// look at the rows, however many near-ties a range holds.
#pragma once
#include "qre_common.cuh"
#include "qre_ptx.cuh"

namespace qre {

struct TopK {              // per-thread candidate state; lives in local memory (rare path only)
  float v[kTopK];
  uint32_t r[kTopK];
  uint32_t chunk;          // head of this (query, range)'s spill chain
  uint32_t fill;           // entries used in the head chunk
};

__device__ __forceinline__ float gate_of(const TopK* tk, float floor_v) {
  return fminf(tk->v[kTopK - 1], fmaxf(floor_v, tk->v[0] - 2.f * kApproxEps));
}

// Chain one candidate.  A new chunk costs one atomic on the batch's bump counter.
static __device__ __noinline__ void spill(TopK* tk, float v, uint32_t row, const GemmArgs* a) {
  uint32_t c = tk->chunk;
  if (c == kChunkLost) return;
  if (c == kNoChunk || tk->fill == kSpillPer) {
    const uint32_t nc = atomicAdd(a->spill_next, 1u);
    if (nc >= a->spill_cap) {          // pool exhausted: the re-score pass settles this query by brute force
      tk->chunk = kChunkLost;
      return;
    }
    if (c != kNoChunk) a->spill_pool[c].n = kSpillPer;
    a->spill_pool[nc].next = c;
    tk->chunk = c = nc;
    tk->fill = 0;
  }
  a->spill_pool[c].e[tk->fill] = PartialEntry{v, row};
  ++tk->fill;
}

// Candidate path: mask (row bound, context, TTL — src/cache.py:274-276) then insert or chain.
// Returns the new gate.  Deliberately not inlined: the fast path must stay tiny.
static __device__ __noinline__ float consider(TopK* tk, float v, long long row, uint64_t myctx, const GemmArgs* a) {
  const float floor_v = a->floor_v;
  if (row >= a->n_rows) return gate_of(tk, floor_v);
  const int4 raw = __ldg(reinterpret_cast<const int4*>(a->meta + row));
  const uint64_t ctx = (static_cast<uint64_t>(static_cast<uint32_t>(raw.y)) << 32) | static_cast<uint32_t>(raw.x);
  const long long ts = static_cast<long long>((static_cast<uint64_t>(static_cast<uint32_t>(raw.w)) << 32) |
                                              static_cast<uint32_t>(raw.z));
  if (ctx != myctx || (a->now_us - ts) > a->ttl_us) return gate_of(tk, floor_v);
  if (v > tk->v[kTopK - 1]) {
    const float ev = tk->v[kTopK - 1];
    const uint32_t er = tk->r[kTopK - 1];
    int i = kTopK - 1;
    while (i > 0 && v > tk->v[i - 1]) {      // strict: equal scores keep the earlier row first
      tk->v[i] = tk->v[i - 1];
      tk->r[i] = tk->r[i - 1];
      --i;
    }
    tk->v[i] = v;
    tk->r[i] = static_cast<uint32_t>(row);
    // the entry pushed off the list stays a contender while it is inside the window of the best
    if (er != kNoRow && ev >= fmaxf(floor_v, tk->v[0] - 2.f * kApproxEps)) spill(tk, ev, er, a);
  } else if (v >= fmaxf(floor_v, tk->v[0] - 2.f * kApproxEps)) {
    spill(tk, v, static_cast<uint32_t>(row), a);
  }
  return gate_of(tk, floor_v);
}

__device__ __forceinline__ void topk_reset(TopK& tk, float floor_v) {
#pragma unroll
  for (int i = 0; i < kTopK; ++i) { tk.v[i] = floor_v; tk.r[i] = kNoRow; }
  tk.chunk = kNoChunk;
  tk.fill = 0;
}

__device__ __forceinline__ void topk_store(const TopK& tk, const GemmArgs& a, long long unit) {
  PartialEntry* out = a.partial + unit * kTopK;
#pragma unroll
  for (int i = 0; i < kTopK; ++i) out[i] = PartialEntry{tk.v[i], tk.r[i]};
  const uint32_t c = tk.chunk;
  if (c != kNoChunk && c != kChunkLost) a.spill_pool[c].n = tk.fill;
  a.spill_head[unit] = c;
}

// Scan one accumulator tile (this thread's lane, kBN columns starting at `taddr`); returns the
// updated gate.  kDump additionally writes the raw scores out (bring-up / tests).
template <bool kDump>
__device__ __forceinline__ float scan_tile(uint32_t taddr, long long row_base, int q, bool active, float gate,
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
    if (mx >= gate) {
#pragma unroll
      for (int i = 0; i < 64; ++i) {
        if (v[i] >= gate) gate = consider(&tk, v[i], row_base + c * 64 + i, myctx, &a);
      }
    }
  }
  return gate;
}

}  // namespace qre
