// Host side of the rollout loop with a host in it (the loop body of src/my_trainers/base_KA_DDPG.py:175-219 replayed as
// one captured CUDA graph per policy step, rollout.py): launch the instantiated graph and wait for the step's results
// WITHOUT cudaDeviceSynchronize — the last CTA of the env kernel stores the step number to a pinned host word after a
// system-scope fence (env_step.cu), and the host spins on that word.  Two driver calls (launch, synchronise) and the
// interpreter between them become one C call.
#include <chrono>
#if defined(__x86_64__) || defined(_M_X64)
#include <immintrin.h>
#define OIS_CPU_RELAX() _mm_pause()
#else
#define OIS_CPU_RELAX() do { } while (0)
#endif
#include "ois_common.cuh"

namespace ois {

static int wait_flag(const volatile uint32_t* flag, uint32_t expect, int64_t timeout_ns) {
  const auto t0 = std::chrono::steady_clock::now();
  uint32_t spins = 0;
  while ((int32_t)(*flag - expect) < 0) {
    OIS_CPU_RELAX();
    if ((++spins & 0xFFFu) == 0u) {
      const int64_t dt = std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now() - t0).count();
      if (timeout_ns > 0 && dt > timeout_ns) {
        // not completed in time: surface a sticky CUDA error if there is one, else report the timeout
        cudaError_t e = cudaPeekAtLastError();
        set_error("ois_host_wait: flag %u did not reach %u within %.3f s (%s)", *flag, expect, 1e-9 * (double)timeout_ns,
                  cudaGetErrorString(e));
        return OIS_ERR_CUDA;
      }
    }
  }
  return OIS_OK;
}

}  // namespace ois

using namespace ois;

extern "C" int ois_graph_launch(void* graph_exec, void* stream) {
  OIS_CHECK_ARG(graph_exec, "null graph");
  OIS_CHECK_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
  return OIS_OK;
}

extern "C" int ois_host_wait(const uint32_t* host_flag, uint32_t expect, int64_t timeout_ns) {
  OIS_CHECK_ARG(host_flag, "null flag");
  return wait_flag(host_flag, expect, timeout_ns);
}

extern "C" int ois_graph_launch_wait(void* graph_exec, void* stream, const uint32_t* host_flag, uint32_t expect,
                                     int64_t timeout_ns) {
  OIS_CHECK_ARG(graph_exec && host_flag, "null argument");
  OIS_CHECK_CUDA(cudaGraphLaunch((cudaGraphExec_t)graph_exec, (cudaStream_t)stream));
  return wait_flag(host_flag, expect, timeout_ns);
}
