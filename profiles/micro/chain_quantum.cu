// Is the cost of a dependent kernel in a chain quantised?  A chain of 20 launches of a kernel that spins for X ns
// (programmatic dependent launches in a stream, the same captured in a CUDA graph, and plain stream order), timed
// with events over 20 repetitions: us per launch as a function of X.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o chain_quantum chain_quantum.cu
#include <cuda_runtime.h>
#include <stdio.h>

__global__ void spin(unsigned long long ns, int pdl, int* sink) {
  if (pdl) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  do { asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); } while (t - t0 < ns);
  if (sink && threadIdx.x == 0 && blockIdx.x == 0) sink[0] = (int)t;
}

static void launch(unsigned long long ns, int pdl, int grid, int* sink, cudaStream_t s) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = 128; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, spin, ns, pdl, sink);
}

int main() {
  int* sink; cudaMalloc(&sink, 4);
  cudaStream_t s; cudaStreamCreate(&s);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int chain = 20, reps = 20;
  printf("spin_ns  stream_pdl  graph_pdl  graph_plain  stream_plain   (us per launch, grid 64)\n");
  for (unsigned long long ns = 0; ns <= 6000; ns += 250) {
    float out[4];
    for (int mode = 0; mode < 4; ++mode) {
      const int pdl = mode < 2, graph = mode == 1 || mode == 2;
      cudaGraphExec_t ge = nullptr;
      if (graph) {
        cudaGraph_t g;
        cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < chain; ++i) launch(ns, pdl, 64, sink, s);
        cudaStreamEndCapture(s, &g);
        cudaGraphInstantiate(&ge, g, 0);
        cudaGraphLaunch(ge, s);
      } else {
        for (int i = 0; i < chain; ++i) launch(ns, pdl, 64, sink, s);
      }
      cudaStreamSynchronize(s);
      cudaEventRecord(a, s);
      for (int r = 0; r < reps; ++r) {
        if (graph) cudaGraphLaunch(ge, s);
        else for (int i = 0; i < chain; ++i) launch(ns, pdl, 64, sink, s);
      }
      cudaEventRecord(b, s);
      cudaStreamSynchronize(s);
      float ms; cudaEventElapsedTime(&ms, a, b);
      out[mode] = ms * 1e3f / (chain * reps);
      if (ge) cudaGraphExecDestroy(ge);
    }
    printf("%6llu  %9.2f  %9.2f  %10.2f  %11.2f\n", ns, out[0], out[1], out[2], out[3]);
  }
  return 0;
}
