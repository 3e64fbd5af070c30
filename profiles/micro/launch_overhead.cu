// Launch overhead probe: event-to-event time of a kernel that does (almost) nothing, as a function of its dynamic shared
// memory, TMEM allocation, thread count, and what ran before it (a memset with the default carve-out, or the same kernel).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o launch_overhead launch_overhead.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>
#include <algorithm>
#include <vector>

template <bool kTmem>
__global__ void probe(int* out) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t slot_s;
  uint32_t* slot = &slot_s;
  if (kTmem) {
    if (threadIdx.x < 32) {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(slot)) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(*slot) : "memory");
  }
  if (threadIdx.x == 0 && out) out[blockIdx.x] = 1;
}

template <bool kTmem>
static float run(int grid, int threads, int smem, int prior, int carve, bool pdl, void* flush, size_t flush_bytes, int* out) {
  cudaFuncSetAttribute(probe<kTmem>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(probe<kTmem>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
  std::vector<float> ts;
  for (int it = 0; it < 15; ++it) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaMemsetAsync(flush, it, flush_bytes);
    if (prior == 1) probe<kTmem><<<grid, threads, smem>>>(out);   // same configuration right before the timed launch
    cudaEventRecord(a);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = threads; cfg.dynamicSmemBytes = smem; cfg.stream = 0;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t le = cudaLaunchKernelEx(&cfg, probe<kTmem>, out);
    cudaEventRecord(b);
    cudaError_t se = cudaDeviceSynchronize();
    if (le != cudaSuccess || se != cudaSuccess) { printf("error: launch %s sync %s\n", cudaGetErrorString(le), cudaGetErrorString(se)); exit(1); }
    float ms; cudaEventElapsedTime(&ms, a, b);
    ts.push_back(ms * 1e3f);
    cudaEventDestroy(a); cudaEventDestroy(b);
  }
  std::sort(ts.begin(), ts.end());
  return ts[ts.size() / 2];
}

int main() {
  void* flush; size_t fb = 256u << 20;
  cudaMalloc(&flush, fb);
  int* out; cudaMalloc(&out, 4096);
  const int smems[] = {0, 49152, 102400, 167936, 230000};
  printf("grid threads smem tmem prior carve pdl  median_us\n");
  for (int smem : smems)
    for (int tm = 0; tm < 2; ++tm)
      for (int prior = 0; prior < 2; ++prior) {
        const int carve = cudaSharedmemCarveoutDefault;
        float t = tm ? run<true>(64, 640, smem, prior, carve, true, flush, fb, out) : run<false>(64, 640, smem, prior, carve, true, flush, fb, out);
        printf("64 640 %6d %d %d default 1 %8.2f\n", smem, tm, prior, t);
      }
  for (int threads : {128, 480, 640, 1024}) {
    float t = run<true>(64, threads, 230000, 0, cudaSharedmemCarveoutDefault, true, flush, fb, out);
    printf("64 %4d 230000 1 0 default 1 %8.2f\n", threads, t);
  }
  for (int grid : {1, 32, 64, 148}) {
    float t = run<true>(grid, 640, 230000, 0, cudaSharedmemCarveoutDefault, true, flush, fb, out);
    printf("%3d 640 230000 1 0 default 1 %8.2f\n", grid, t);
  }
  float t = run<true>(64, 640, 230000, 0, cudaSharedmemCarveoutDefault, false, flush, fb, out);
  printf("64 640 230000 1 0 default 0(no pdl) %8.2f\n", t);
  t = run<false>(64, 640, 0, 0, cudaSharedmemCarveoutMaxShared, true, flush, fb, out);
  printf("64 640 0 0 0 maxshared 1 %8.2f\n", t);
  return 0;
}
