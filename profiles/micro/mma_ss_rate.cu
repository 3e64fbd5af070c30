// Diagnostics: cycles per tcgen05.mma (kind::f16, bf16 -> fp32, cta_group::1, M = 128, K = 16) with BOTH operands in shared
// memory in the canonical no-swizzle layouts mlp_tct.cu uses: A K-major (LBO 2048, SBO 128, a fresh 4 KB slice per MMA as in
// a weight tile), B either K-major or MN-major (LBO 128, SBO 8192: the env-major activations).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../options_in_simulation_b200/csrc -o mma_ss_rate mma_ss_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace ois;

template <int N, bool B_MN, bool A_TMEM>
__global__ void __launch_bounds__(128, 1) rate_kernel(int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];   // [0, 128 KB): four 32 KB "weight tiles"; [128 KB, 192 KB): activations
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 196608 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 0) ptx::tmem_alloc<512>(&slot);
  ptx::fence_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 1) {
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::idesc_bf16(128, N) | (B_MN ? (1u << 16) : 0u);
    constexpr uint32_t hi_a = ptx::smem_desc_hi(128);
    const uint32_t hi_b = B_MN ? ptx::smem_desc_hi(8192) : ptx::smem_desc_hi(128);
    const uint32_t a0 = ptx::smem_u32(smem), b0 = ptx::smem_u32(smem) + 131072;
    const long long t0 = clock64();
    for (int it = 0; it < reps / 32; ++it) {
      if (leader) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {   // 4 tiles x 8 K steps, four accumulators (as layer 2: mb = j / 8)
          const uint32_t d = tmem + 256 + (uint32_t)(j >> 3) * (uint32_t)N % 256;
          const uint64_t bd = B_MN ? ptx::make_desc(ptx::smem_desc_lo(b0 + (j & 7) * 256, 128), hi_b)
                                   : ptx::make_desc(ptx::smem_desc_lo(b0 + (j & 7) * (N * 32), N * 16), hi_b);
          if (A_TMEM) ptx::mma_ts(d, tmem + (j & 7) * 8, bd, idesc, true);
          else ptx::mma_ss(d, ptx::make_desc(ptx::smem_desc_lo(a0 + (j >> 3) * 32768 + (j & 7) * 4096, 2048), hi_a), bd, idesc, true);
        }
      }
      __syncwarp();
    }
    if (leader) ptx::mma_commit(&bar);
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && leader) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

template <int N, bool B_MN, bool A_TMEM>
static void run(long long* d) {
  const int reps = 4096;
  cudaFuncSetAttribute(rate_kernel<N, B_MN, A_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 196608);
  rate_kernel<N, B_MN, A_TMEM><<<148, 128, 196608>>>(reps, d);
  long long c = 0;
  cudaError_t e = cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); exit(1); }
  printf("M=128 N=%3d K=16  A %s  B %s: %6.1f cycles / MMA (%5.0f MAC/clk/SM)\n", N, A_TMEM ? "TMEM        " : "smem K-major",
         B_MN ? "smem MN-major" : "smem K-major ", (double)c / reps, 128.0 * N * 16 * reps / (double)c);
}

int main() {
  long long* d;
  cudaMalloc(&d, 8);
  run<32, false, false>(d); run<32, true, false>(d);
  run<64, false, false>(d); run<64, true, false>(d);
  run<128, false, false>(d); run<128, true, false>(d);
  run<256, false, false>(d); run<256, true, false>(d);
  run<64, false, true>(d); run<64, true, true>(d);
  run<128, false, true>(d); run<256, false, true>(d);
  return 0;
}
