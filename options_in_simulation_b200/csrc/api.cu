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
