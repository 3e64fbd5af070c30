// Host-in-the-loop latency on an idle GPU: wall time from the launch call to a completion word in pinned host memory,
// for two dependent kernels (spin A ns, then spin B ns and write the word) issued (a) as one cudaGraphLaunch, (b) as two
// eager programmatic-dependent launches, (c) as two eager plain launches; and for ONE kernel of the same total length (the
// fused policy step) issued (d) as a one-node graph, (e) as one eager launch with the PDL attribute, (f) as one plain launch.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o idle_launch idle_launch.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <chrono>
#include <algorithm>
#include <vector>

__global__ void spin(unsigned long long ns, int pdl, volatile unsigned* flag, unsigned value) {
  if (pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns);
  if (flag && threadIdx.x == 0 && blockIdx.x == 0) { __threadfence_system(); *flag = value; }
}

static void launch(unsigned long long ns, int pdl, volatile unsigned* flag, unsigned value, cudaStream_t s) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = 64; cfg.blockDim = 128; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, spin, ns, pdl, flag, value);
}

int main() {
  unsigned* flag; cudaHostAlloc(&flag, 4, cudaHostAllocMapped);
  *flag = 0;
  cudaStream_t s; cudaStreamCreate(&s);
  const unsigned long long A = 15000, B = 5000;
  printf("kernels: %llu + %llu ns of spinning; wall us from the launch call to the host seeing the completion word\n", A, B);
  for (int mode = 0; mode < 6; ++mode) {
    std::vector<double> ts;
    unsigned v = 0;
    for (int it = 0; it < 300; ++it) {
      ++v;
      cudaGraphExec_t ge = nullptr;
      if (mode == 3) {
        cudaGraph_t g;
        cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        launch(A + B, 1, flag, 1, s);
        cudaStreamEndCapture(s, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        cudaGraphLaunch(ge, s);
        cudaStreamSynchronize(s);
      }
      if (mode == 0) {   // the word's value is a kernel argument: one graph per value would be unfair, so the graph writes `it`-independent 1
        cudaGraph_t g;
        cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        launch(A, 1, nullptr, 0, s);
        launch(B, 1, flag, 1, s);
        cudaStreamEndCapture(s, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        cudaGraphLaunch(ge, s);      // warm the exec
        cudaStreamSynchronize(s);
      }
      *flag = 0;
      cudaStreamSynchronize(s);
      auto t0 = std::chrono::steady_clock::now();
      if (mode == 0 || mode == 3) cudaGraphLaunch(ge, s);
      else if (mode >= 4) launch(A + B, mode == 4, flag, 1, s);
      else { launch(A, mode == 1, nullptr, 0, s); launch(B, mode == 1, flag, 1, s); }
      while (*(volatile unsigned*)flag == 0) { }
      auto t1 = std::chrono::steady_clock::now();
      ts.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
      cudaStreamSynchronize(s);
      if (ge) cudaGraphExecDestroy(ge);
    }
    std::sort(ts.begin(), ts.end());
    printf("%-28s median %6.1f us  min %6.1f us  (kernels alone: %.1f us)\n",
           mode == 0 ? "one graph launch" : mode == 1 ? "two eager PDL launches" : mode == 2 ? "two eager plain launches" :
           mode == 3 ? "ONE kernel: one-node graph" : mode == 4 ? "ONE kernel: eager PDL" : "ONE kernel: eager plain", ts[150], ts[0], (A + B) * 1e-3);
  }
  return 0;
}
