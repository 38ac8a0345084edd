// Shared declarations of the on-GPU query embedder (include/qre_encoder_b200.h).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/qre_encoder_b200.h"

namespace qre {

// Tile of the encoder GEMM: 128 tokens (one TMEM lane each) x 192 output features.  192 divides
// every feature count of the MiniLM / BERT-base family (384, 768, 1152, 1536, 2304, 3072), and
// two fp32 accumulators of 192 columns fit the 512 TMEM columns for epilogue/MMA overlap.
constexpr int kEncBM = 128;
constexpr int kEncBN = 192;
constexpr int kEncBK = 64;
constexpr int kEncStages = 6;
constexpr int kEncRowPad = 256;    // activation buffers are padded to whole CTA-pair tiles (2 x 128 tokens)

enum EncEpilogue { kEpiBias = 0, kEpiGelu = 1, kEpiResidual = 2, kEpiResidualLN = 3 };

struct EncGemmArgs {
  int num_kb;                    // input features / 64
  int tiles_m, tiles_n;          // ceil(tokens / 128), ceil(features / 192)
  const float* bias;             // [tiles_n * 192], zero padded
  // kEpiResidualLN only (features == 384 = two feature tiles): out = LayerNorm(acc + bias + residual)
  const float* ln_g = nullptr;   // [384]
  const float* ln_b = nullptr;   // [384]
  float ln_eps = 0.f;
};
// Output / residual boxes of the GEMM epilogue: 32 tokens x 64 features, 128-byte swizzle.
constexpr int kEncBoxRows = 32;
constexpr int kEncBoxCols = 64;

cudaError_t enc_gemm_init();
// Returns the number of kernel launches (1).
// tmap_out / tmap_res: [tokens padded to 128, features] bf16 with the kEncBox box; tmap_res may be null
// unless epilogue == kEpiResidual.  Rows past the batch's tokens are computed and stored like any
// other (the buffers are padded); features past the tensor's width are clipped by the TMA unit.
int launch_enc_gemm(const CUtensorMap& tmap_a, const CUtensorMap& tmap_w, const CUtensorMap& tmap_out,
                    const CUtensorMap* tmap_res, const EncGemmArgs& a, int epilogue, int num_sms, cudaStream_t st,
                    bool after_kernel = false);

// ---- programmatic dependent launch ------------------------------------------------------------
// The encoder is a chain of ~45 short kernels on one stream.  Each is launched with the
// "programmatic stream serialization" attribute: its CTAs may be scheduled (and run their set-up:
// barrier init, TMEM allocation, descriptor prefetch) while the previous kernel is still draining,
// and wait in pdl_wait() until that kernel has completed and its writes are visible.  Rules every
// kernel of the chain follows: pdl_launch_dependents() first, then pdl_wait() before the first global
// read OR write — every link waits, so completion is transitive along the chain.  Without the launch
// attribute both instructions do nothing.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_chain(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                bool after_kernel, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = after_kernel ? 1 : 0;      // the first kernel of a chain follows copies / unrelated work: plain launch
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// Sets the thread's last-error string (qre_last_error) and returns `code`.
int set_error(int code, const char* fmt, ...);
// bf16 2-D K-major tensor map with the 128-byte swizzle the UMMA descriptors assume.
int make_tensor_map_bf16(CUtensorMap* map, void* base, uint64_t inner, uint64_t rows, uint32_t box_inner,
                         uint32_t box_rows);

}  // namespace qre
