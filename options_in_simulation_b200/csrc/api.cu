// Error plumbing and build identification for libois_b200.so.
#include <stdarg.h>
#include <string.h>
#include "ois_common.cuh"

namespace ois {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace ois

extern "C" const char* ois_last_error_string(void) { return ois::g_err; }
extern "C" int ois_abi_version(void) { return OIS_ABI_VERSION; }
extern "C" const char* ois_build_info(void) { return "sm_100a; env: fused fp32 SIMT; mlp: tcgen05+TMA bf16 / fp32 SIMT"; }
extern "C" int ois_struct_size(int which) {
  switch (which) {
    case 0: return (int)sizeof(OisEnvParams);
    case 1: return (int)sizeof(OisEnvState);
    case 2: return (int)sizeof(OisEnvIO);
    case 3: return (int)sizeof(OisMlpShape);
    case 4: return (int)sizeof(OisScaler);
    case 5: return (int)sizeof(OisActIO);
    case 6: return (int)sizeof(OisUpdateIO);
    case 7: return (int)sizeof(OisStepCtl);
    case 8: return (int)sizeof(OisSampleCtl);
    default: return -1;
  }
}

// Stand-alone advance of the step-control block (ois_common.cuh: step_ctl_advance) for callers that do not let the env
// kernel do it (OisStepCtl.self_advance = 0): a single thread, launched last in the captured step.
namespace ois {
__global__ void step_ctl_advance_kernel(OisStepCtl* c) { step_ctl_advance(c); }
}

extern "C" int ois_step_ctl_advance(OisStepCtl* ctl, void* stream) {
  OIS_CHECK_ARG(ctl, "null ctl");
  ois::step_ctl_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(ctl);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

namespace ois {
int& launch_budget() {
  static int budget = -1;
  return budget;
}
}  // namespace ois

// diagnostics hook (deliberately not declared in include/ois_b200.h): see launch_budget() in ois_common.cuh
extern "C" int ois_debug_launch_budget(int n) {
  ois::launch_budget() = n;
  return 0;
}
