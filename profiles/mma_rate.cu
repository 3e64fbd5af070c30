// Diagnostics: issue-to-completion rate of tcgen05.mma (kind::f16, bf16 -> fp32, cta_group::1, M = 128) for the shapes
// the actor kernel uses, with A from shared memory and A from tensor memory.  One CTA per SM, one issuing thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../options_in_simulation_b200/csrc -o mma_rate mma_rate.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"

using namespace ois;

template <int N, int NACC, bool FROM_TMEM>
__global__ void __launch_bounds__(128, 1) rate_kernel(int reps, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 65536 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 0) ptx::tmem_alloc<512>(&slot);
  ptx::fence_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 1) {      // whole warp runs the control flow, one elected lane issues (as the actor kernel does)
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::idesc_bf16(128, N);
    constexpr uint32_t hi = ptx::smem_desc_hi(128);
    const uint32_t a_lo = ptx::smem_desc_lo(ptx::smem_u32(smem), 2048);
    const uint32_t b_lo = ptx::smem_desc_lo(ptx::smem_u32(smem) + 8192, 16 * N);
    if (leader) {
      for (int i = 0; i < 8; ++i) ptx::mma_ss(tmem + 256, ptx::make_desc(a_lo, hi), ptx::make_desc(b_lo, hi), idesc, i != 0);
      ptx::mma_commit(&bar);
    }
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    const long long t0 = clock64();
    for (int it = 0; it < reps / 16; ++it) {
      if (leader) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t d = tmem + 256 + (uint32_t)(j % NACC) * (uint32_t)N;   // NACC independent accumulators, round robin
          if (FROM_TMEM) ptx::mma_ts(d, tmem + (j & 7) * 8, ptx::make_desc(b_lo + (j & 3) * 16, hi), idesc, true);
          else ptx::mma_ss(d, ptx::make_desc(a_lo + (j & 1) * 16, hi), ptx::make_desc(b_lo + (j & 3) * 16, hi), idesc, true);
        }
      }
      __syncwarp();
    }
    if (leader) ptx::mma_commit(&bar);
    __syncwarp();
    ptx::mbar_wait(&bar, 1);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && leader) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

// Interference test: warp 1 issues N=64 A-from-TMEM MMAs (the layer-2 shape) while, optionally, four threads stream
// 16 KB bulk copies into shared memory and / or eight warps tcgen05.ld 64 columns in a loop (the epilogue pattern).
__global__ void __launch_bounds__(448, 1) mix_kernel(const uint8_t* img, int reps, int with_bulk, int with_ldtm, int bstride, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, full[4][4];
  __shared__ uint32_t slot;
  __shared__ volatile int stop;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar, 1);
    for (int w = 0; w < 4; ++w) for (int i = 0; i < 4; ++i) ptx::mbar_init(&full[w][i], 1);
    ptx::fence_barrier_init();
    stop = 0;
  }
  if (warp == 0) ptx::tmem_alloc<512>(&slot);
  ptx::fence_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (warp == 1) {
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::idesc_bf16(128, 64);
    constexpr uint32_t hi = ptx::smem_desc_hi(128);
    const uint32_t b_lo = ptx::smem_desc_lo(ptx::smem_u32(smem), 1024);
    const long long t0 = clock64();
    for (int it = 0; it < reps / 16; ++it) {
      if (leader) {
#pragma unroll
        for (int j = 0; j < 16; ++j)
          ptx::mma_ts(tmem + 256 + (j & 1) * 64, tmem + (j & 7) * 8 + (j >> 3) * 64, ptx::make_desc(b_lo + j * bstride, hi), idesc, true);
      }
      __syncwarp();
    }
    if (leader) ptx::mma_commit(&bar);
    __syncwarp();
    ptx::mbar_wait(&bar, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && leader) out[0] = t1 - t0;
    stop = 1;
  } else if (warp >= 2 && warp < 10) {
    if (with_ldtm) {
      uint32_t acc = 0;
      const uint32_t t_lane = tmem + ((uint32_t)((warp & 3) * 32) << 16);
      while (!stop) {
        uint32_t v[32];
        ptx::tmem_ld32(t_lane + 384, v);
        ptx::tmem_ld_wait();
        acc += v[lane & 31];
        ptx::tmem_ld32(t_lane + 416, v);
        ptx::tmem_ld_wait();
        acc += v[lane & 31];
      }
      if (acc == 0x12345678u) out[1] = acc;
    }
  } else if (warp >= 10) {
    if (with_bulk && lane == 0) {
      const int w = warp - 10;
      uint8_t* base = smem + 32768 + (size_t)w * 2 * 16384;
      for (int i = 0; !stop; ++i) {
        if (i >= 2) ptx::mbar_wait(&full[w][i & 1], ((i - 2) >> 1) & 1);
        ptx::mbar_expect_tx(&full[w][i & 1], 16384);
        ptx::bulk_g2s(base + (i & 1) * 16384, img + (size_t)((i * 4 + w) % 40) * 16384, 16384, &full[w][i & 1]);
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

// Queue-depth test: between groups of `group` MMAs the issuing warp does what the actor kernel does per weight stage
// (a satisfied mbarrier wait + tcgen05.fence::after_thread_sync + __syncwarp + a tcgen05.commit).  If the tensor pipe
// keeps its 32 cycles / MMA the issue queue absorbs that overhead; otherwise the overhead is exposed.
__global__ void __launch_bounds__(128, 1) gap_kernel(int reps, int group, int overhead, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar, done_bar, junk[4];
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar, 1);
    ptx::mbar_init(&done_bar, 1);
    for (int i = 0; i < 4; ++i) ptx::mbar_init(&junk[i], 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) ptx::tmem_alloc<512>(&slot);
  ptx::fence_async_smem();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) ptx::mbar_arrive(&bar);      // phase 0 of `bar` is complete: waits on parity 0 succeed at once
  __syncthreads();
  if (warp == 1) {
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::idesc_bf16(128, 64);
    constexpr uint32_t hi = ptx::smem_desc_hi(128);
    const uint32_t b_lo = ptx::smem_desc_lo(ptx::smem_u32(smem) + 8192, 1024);
    const long long t0 = clock64();
    for (int it = 0; it < reps / group; ++it) {
      for (int o = 0; o < overhead; ++o) {
        ptx::mbar_wait(&bar, 0);
        ptx::tc_fence_after();
      }
      if (leader) {
        for (int j = 0; j < group; ++j)
          ptx::mma_ts(tmem + 256 + (j & 1) * 64, tmem + (j & 7) * 8, ptx::make_desc(b_lo + (j & 3) * 16, hi), idesc, true);
      }
      __syncwarp();
      if (overhead) {
        if (leader) ptx::mma_commit(&junk[it & 3]);
        __syncwarp();
      }
    }
    if (leader) ptx::mma_commit(&done_bar);
    __syncwarp();
    ptx::mbar_wait(&done_bar, 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && leader) out[0] = t1 - t0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc<512>(tmem);
}

template <int N, int NACC, bool FROM_TMEM>
void run(long long* d) {
  const int reps = 2048;
  cudaFuncSetAttribute(rate_kernel<N, NACC, FROM_TMEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  rate_kernel<N, NACC, FROM_TMEM><<<148, 128, 65536>>>(reps, d);
  long long c = 0;
  cudaError_t e = cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); exit(1); }
  printf("acc %d  A from %s  M=128 N=%3d K=16: %7.1f cycles / MMA  (%.0f MAC/clk/SM)\n", NACC, FROM_TMEM ? "TMEM" : "smem", N,
         (double)c / reps, 128.0 * N * 16 * reps / (double)c);
}

// L2 -> shared-memory streaming rate of 1-D bulk copies (the weight ring of the actor kernel): one thread keeps `depth`
// copies of `bytes` in flight from a 657 KB L2-resident image.
__global__ void __launch_bounds__(128, 1) bulk_kernel(const uint8_t* img, int bytes, int depth, int reps, int issuers, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[4][16];
  if (threadIdx.x == 0) {
    for (int w = 0; w < 4; ++w) for (int i = 0; i < 16; ++i) ptx::mbar_init(&full[w][i], 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  long long t0 = clock64();
  if ((threadIdx.x & 31) == 0 && w < issuers) {   // `issuers` threads (one per warp), each with its own ring region
    const int stages = 655360 / bytes;
    uint8_t* base = smem + (size_t)w * depth * bytes;
    for (int i = 0; i < reps + depth; ++i) {
      if (i >= depth) ptx::mbar_wait(&full[w][(i - depth) % depth], ((i - depth) / depth) & 1);
      if (i < reps) {
        const int slot = i % depth;
        ptx::mbar_expect_tx(&full[w][slot], bytes);
        ptx::bulk_g2s(base + slot * bytes, img + (size_t)((i * issuers + w) % stages) * bytes, bytes, &full[w][slot]);
      }
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
}

int main() {
  {
    long long* d;
    uint8_t* img;
    cudaMalloc(&d, 8);
    cudaMalloc(&img, 655360);
    cudaMemset(img, 1, 655360);
    cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 196608);
    for (int issuers : {1, 2, 4})
      for (int bytes : {4096, 16384, 32768})
        for (int depth : {1, 2, 4}) {
          if (bytes * depth * issuers > 196608) continue;
          const int reps = 1000;
          bulk_kernel<<<148, 128, 196608>>>(img, bytes, depth, reps, issuers, d);
          long long c = 0;
          if (cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("bulk error\n"); return 1; }
          printf("bulk issuers %d  %5d B x depth %2d: %6.1f B/clk/SM  (%.0f GB/s per SM at 1.965 GHz), %6.0f cycles per copy\n", issuers, bytes, depth,
                 (double)bytes * reps * issuers / c, (double)bytes * reps * issuers / c * 1.965, (double)c / reps);
        }
  }
  {
    long long* d;
    uint8_t* img;
    cudaMalloc(&d, 16);
    cudaMalloc(&img, 655360);
    cudaMemset(img, 1, 655360);
    cudaFuncSetAttribute(mix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 131072);
    for (int bstride : {1, 128})
      for (int bulk = 0; bulk < 2; ++bulk)
        for (int ldtm = 0; ldtm < 2; ++ldtm) {
          mix_kernel<<<148, 448, 32768 + 131072>>>(img, 4096, bulk, ldtm, bstride, d);
          long long c = 0;
          if (cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("mix error\n"); return 1; }
          printf("mix  N=64 A-from-TMEM MMAs, B tiles %4d B apart, bulk copies %d, tcgen05.ld loops %d: %6.1f cycles / MMA\n", bstride * 16, bulk,
                 ldtm, (double)c / 4096);
        }
  }
  {
    long long* d;
    cudaMalloc(&d, 16);
    cudaFuncSetAttribute(gap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    for (int group : {8, 32})
      for (int overhead : {0, 1, 2, 4}) {
        gap_kernel<<<148, 128, 32768>>>(4096, group, overhead, d);
        long long c = 0;
        if (cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("gap error\n"); return 1; }
        printf("gap  groups of %2d MMAs, %d x (mbarrier wait + fence) + commit between groups: %6.1f cycles / MMA\n", group, overhead,
               (double)c / 4096);
      }
  }
  long long* d;
  cudaMalloc(&d, 8);
  run<16, 1, false>(d); run<64, 1, false>(d); run<128, 1, false>(d); run<256, 1, false>(d);
  run<16, 1, true>(d); run<64, 1, true>(d); run<128, 1, true>(d); run<256, 1, true>(d);
  run<64, 2, true>(d); run<64, 4, true>(d); run<128, 2, true>(d); run<64, 2, false>(d); run<64, 4, false>(d); run<128, 2, false>(d);
  return 0;
}
