// Microbenchmark: latency and per-SM throughput of mma.sync.m16n8k16 (bf16 -> fp32) on this GPU, plus
// ldmatrix and shuffle latency.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/mma_lat
// tools/micro/mma_sync_latency.cu ; run on the GPU box.  Informs the attention kernels (qre_encoder.cu).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

__device__ __forceinline__ void mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int kChains>
__global__ void mma_kernel(float* out, long long* cycles, int iters) {
  float c[kChains][4];
  for (int i = 0; i < kChains; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.f;
  const uint32_t a = 0x3f803f80u + threadIdx.x, b = 0x3f803f80u;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int i = 0; i < kChains; ++i) mma(c[i], a, a, a, a, b, b);
  const long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < kChains; ++i) s += c[i][0] + c[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
}

__global__ void shfl_kernel(float* out, long long* cycles, int iters) {
  float v = threadIdx.x;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) v = __shfl_xor_sync(0xffffffffu, v, 1) + 1.f;
  const long long t1 = clock64();
  out[threadIdx.x] = v;
  if (threadIdx.x == 0) *cycles = t1 - t0;
}

__global__ void ldmatrix_kernel(uint32_t* out, long long* cycles, int iters) {
  __shared__ __align__(16) uint16_t sm[64 * 40];
  for (int i = threadIdx.x; i < 64 * 40; i += 32) sm[i] = 8 * (i % 7);      // every loaded word is a small row offset
  __syncwarp();
  uint32_t addr = static_cast<uint32_t>(__cvta_generic_to_shared(sm + (threadIdx.x & 7) * 40 + (threadIdx.x >> 3) * 8));
  uint32_t r0 = 0, r1, r2, r3;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr + ((r0 & 0x30))));
  }
  const long long t1 = clock64();
  out[threadIdx.x] = r0 + r1 + r2 + r3;
  if (threadIdx.x == 0) *cycles = t1 - t0;
}

int main() {
  float* out;
  long long *cyc, h;
  cudaMalloc(&out, 1 << 22);
  cudaMalloc(&cyc, 8);
  const int iters = 4096;
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  printf("{\"gpu\": \"%s\"", p.name);
#define RUN(name, launch, per)                                    \
  launch;                                                         \
  launch;                                                         \
  cudaDeviceSynchronize();                                        \
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);                 \
  printf(", \"%s\": %.2f", name, (double)h / iters / (per));
  RUN("mma_dependent_latency_cycles", (mma_kernel<1><<<1, 32>>>(out, cyc, iters)), 1)
  RUN("mma_1warp_8chains_cycles_per_mma", (mma_kernel<8><<<1, 32>>>(out, cyc, iters)), 8)
  RUN("mma_4warps_8chains_cycles_per_mma_per_sm", (mma_kernel<8><<<1, 128>>>(out, cyc, iters)), 32)
  RUN("mma_16warps_8chains_cycles_per_mma_per_sm", (mma_kernel<8><<<1, 512>>>(out, cyc, iters)), 128)
  RUN("mma_32warps_4chains_cycles_per_mma_per_sm", (mma_kernel<4><<<1, 1024>>>(out, cyc, iters)), 128)
  RUN("shfl_dependent_latency_cycles", (shfl_kernel<<<1, 32>>>(out, cyc, iters)), 1)
  RUN("ldmatrix_x4_dependent_latency_cycles", (ldmatrix_kernel<<<1, 32>>>((uint32_t*)out, cyc, iters)), 1)
  printf(", \"cuda_error\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
