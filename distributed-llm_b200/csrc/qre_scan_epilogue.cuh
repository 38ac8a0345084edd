// qre_scan_epilogue.cuh — the fused top-k epilogue shared by both scan kernels
// (qre_gemm_topk.cu single-CTA, qre_gemm_topk2.cu CTA-pair).
//
// One epilogue thread owns one query (one TMEM lane).  Per 256-column accumulator tile it reads
// 4 chunks of 64 scores with tcgen05.ld, reduces each chunk to its maximum with 3-input FMNMX
// and compares it with the query's current gate; only chunks that can still matter take the
// candidate path (append to the (query, range)'s list).
//
// Candidate path.  The exact winner of a query has approx >= a* - 2 eps where a*
// is the best approximate score over all VALID rows (same context, not expired); nothing under the
// hit floor can hit.  The scan does not look at a row's (context, timestamp) itself: a global
// load per candidate in the thread that also has to hand the accumulator back stalls the tile
// pipeline of the whole CTA pair (measured: a cache of paraphrase clusters ran at half speed).
// Instead every score >= gate is APPENDED, unvalidated, with plain stores: the first 8 of a
// (query, row range) into its slots of `partial`, the rest into a chain of 128-byte spill chunks
// (15 candidates + link) taken from a per-batch pool in groups of 1, 2, 4, 8, 8, ... chunks per
// atomic.  The gate starts at the floor.  Only when a thread has to take a new group (>= 8
// candidates in one range already) does it validate a SAMPLE — the inline slots the first time,
// afterwards the chunk it has just filled — with independent loads under the atomic's latency,
// and lifts the gate to max(floor, best valid sample - 2 eps).  A sample's best is <= a*, so
// nothing with approx >= a* - 2 eps is ever dropped, however few rows were validated; the
// re-score kernel applies the row mask to everything that was appended and takes a* from there.
#pragma once
#include "qre_common.cuh"
#include "qre_ptx.cuh"

namespace qre {

// Append cursor of one (query, row range): lives in the epilogue thread's registers.
struct Cursor {
  PartialEntry* nxt;       // where the next candidate goes
  uint32_t room;           // entries left at nxt (inline slots first, then the head spill chunk)
  float gate;
};

struct TopK {              // the rest of the per-thread candidate state; local memory, touched only at chunk boundaries
  PartialEntry* inl;       // this (query, range)'s kTopK inline slots in `partial`
  uint32_t chunk;          // head of the spill chain: kNoChunk / kChunkLost / pool index
  uint32_t free_next;      // next unused chunk of the group reserved last
  uint32_t free_left;      // chunks left in that group
  uint32_t group;          // size of the next group to reserve
};

constexpr uint32_t kSpillGroupMax = 8;

__device__ __forceinline__ bool row_valid(const RowMeta* meta, uint32_t row, uint64_t myctx, long long now_us,
                                          long long ttl_us) {
  const int4 raw = __ldg(reinterpret_cast<const int4*>(meta + row));
  const uint64_t ctx = (static_cast<uint64_t>(static_cast<uint32_t>(raw.y)) << 32) | static_cast<uint32_t>(raw.x);
  const long long ts = static_cast<long long>((static_cast<uint64_t>(static_cast<uint32_t>(raw.w)) << 32) |
                                              static_cast<uint32_t>(raw.z));
  return ctx == myctx && (now_us - ts) <= ttl_us;     // src/cache.py:274-276, 523-524
}

// Best approximate score among the VALID entries of e[0..n) (n <= kSpillPer); -inf if none.
// Entries were written by this thread (program order makes them visible to it).
__device__ __forceinline__ float best_valid_of(const PartialEntry* e, uint32_t n, uint64_t myctx, const GemmArgs* a) {
  PartialEntry c[kSpillPer];
#pragma unroll
  for (int i = 0; i < kSpillPer; ++i) {
    c[i].row = kNoRow;
    if (i < static_cast<int>(n)) c[i] = e[i];
  }
  bool ok[kSpillPer];
#pragma unroll
  for (int i = 0; i < kSpillPer; ++i)
    ok[i] = c[i].row != kNoRow && row_valid(a->meta, c[i].row, myctx, a->now_us, a->ttl_us);
  float best = -INFINITY;
#pragma unroll
  for (int i = 0; i < kSpillPer; ++i)
    if (ok[i]) best = fmaxf(best, c[i].v);
  return best;
}

// The cursor ran out of room (inline slots or the head chunk are full) and candidate (v, row)
// is waiting: link a new head chunk — reserving the next group of chunks and validating a sample
// when the last group is used up — store the candidate there and return the new cursor.
// Deliberately not inlined: the fast path must stay tiny.
static __device__ __noinline__ Cursor refill(TopK* tk, float gate, float v, uint32_t row, uint64_t myctx,
                                             const GemmArgs* a) {
  uint32_t c = tk->chunk;
  if (c == kChunkLost) return Cursor{nullptr, 0u, gate};
  if (tk->free_left == 0) {
    // reserve the next group; validate a sample while the atomic is in flight
    const uint32_t g = tk->group;
    const uint32_t first = atomicAdd(a->spill_next, g);
    const float bv = (c == kNoChunk) ? best_valid_of(tk->inl, kTopK, myctx, a)
                                     : best_valid_of(a->spill_pool[c].e, kSpillPer, myctx, a);
    gate = fmaxf(gate, bv - 2.f * kApproxEps);
    if (first + g > a->spill_cap || first + g < first) {   // pool exhausted: the re-score pass settles this
      if (c != kNoChunk) a->spill_pool[c].n = kSpillPer;   // query by brute force
      tk->chunk = kChunkLost;
      return Cursor{nullptr, 0u, gate};
    }
    tk->free_next = first;
    tk->free_left = g;
    tk->group = g < kSpillGroupMax ? g * 2 : g;
  }
  const uint32_t nc = tk->free_next;
  ++tk->free_next;
  --tk->free_left;
  if (c != kNoChunk) a->spill_pool[c].n = kSpillPer;
  SpillChunk* ch = a->spill_pool + nc;
  ch->next = c;
  tk->chunk = nc;
  if (v >= gate) {                                         // (the lifted gate may already exclude it)
    ch->e[0] = PartialEntry{v, row};
    return Cursor{ch->e + 1, kSpillPer - 1, gate};
  }
  return Cursor{ch->e, kSpillPer, gate};
}

// One score that passed the gate: append it (registers + one store), or take the slow call.
__device__ __forceinline__ void consider(Cursor& cu, TopK& tk, float v, long long row, uint64_t myctx,
                                         const GemmArgs& a) {
  if (row >= a.n_rows) return;
  if (cu.room != 0) {
    *cu.nxt = PartialEntry{v, static_cast<uint32_t>(row)};
    ++cu.nxt;
    --cu.room;
  } else {
    cu = refill(&tk, cu.gate, v, static_cast<uint32_t>(row), myctx, &a);
  }
}

__device__ __forceinline__ void topk_reset(Cursor& cu, TopK& tk, const GemmArgs& a, long long unit, float gate0) {
  tk.inl = a.partial + unit * kTopK;
  tk.chunk = kNoChunk;
  tk.free_left = 0;
  tk.group = 1;
  cu.nxt = tk.inl;
  cu.room = kTopK;
  cu.gate = gate0;
}

// End of a (query, range): mark the unused inline slots empty and publish the chain head.
__device__ __forceinline__ void topk_store(const Cursor& cu, const TopK& tk, const GemmArgs& a, long long unit) {
  const uint32_t c = tk.chunk;
  const uint32_t n_inl = c == kNoChunk ? kTopK - cu.room : kTopK;
  PartialEntry* out = a.partial + unit * kTopK;
#pragma unroll
  for (int i = 0; i < kTopK; ++i)
    if (i >= static_cast<int>(n_inl)) out[i] = PartialEntry{0.f, kNoRow};
  if (c != kNoChunk && c != kChunkLost) a.spill_pool[c].n = kSpillPer - cu.room;
  a.spill_head[unit] = c;
}

// Scan one accumulator tile (this thread's lane, kBN columns starting at `taddr`); the append
// cursor is updated in place.  kDump additionally writes the raw scores out (bring-up / tests).
template <bool kDump>
__device__ __forceinline__ void scan_tile(uint32_t taddr, long long row_base, int q, bool active, Cursor& cu,
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
    // maxima of the 8 groups of 8 consecutive scores (3-input FMNMX), then of the chunk: a chunk
    // that holds a candidate costs 8 group tests plus 8 score tests per group that has one,
    // instead of 64 score tests for the whole warp
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float* w = v + 8 * j;
      g[j] = fmaxf(fmaxf(fmaxf(fmaxf(w[0], w[1]), w[2]), fmaxf(fmaxf(w[3], w[4]), w[5])), fmaxf(w[6], w[7]));
    }
    const float mx = fmaxf(fmaxf(fmaxf(fmaxf(g[0], g[1]), g[2]), fmaxf(fmaxf(g[3], g[4]), g[5])), fmaxf(g[6], g[7]));
    if (mx >= cu.gate) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (g[j] >= cu.gate) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            if (v[8 * j + i] >= cu.gate) consider(cu, tk, v[8 * j + i], row_base + c * 64 + 8 * j + i, myctx, a);
          }
        }
      }
    }
  }
}

}  // namespace qre
