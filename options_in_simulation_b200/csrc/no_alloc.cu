// ois_mlp_forward_static: the fp32 MLP forward of ois_mlp_forward for callers that may not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// ois_mlp_forward keeps the three hidden activations ([rows, 512], [rows, 512], [rows, 128] fp32) in the workspace that
// ois_agent_create cudaMalloc'ed.  This variant needs no handle: the same scratch lives in ONE file-scope device array
// of fixed size (OIS_STATIC_MLP_MAX_ROWS rows = 18 MiB), which every GPU that loads the library holds until the process
// exits.  The call path below calls no cudaMalloc / cudaFree and no allocating library routine; it launches the same
// four gemm_nt kernels (mlp_fp32.cu) through the same mlp_forward_fp32 (agent.cu) as ois_mlp_forward, so the results are
// bit-identical.
//
// The array is shared by every call on the device and keeps the last call's contents.  Nothing in it has to start at
// zero: each layer writes every hidden element [0, rows) x [0, width) before the next layer reads it, and nothing
// outside that range is read.  Calls that overlap on the GPU (two streams, two host threads) race on it: callers
// serialise them (one stream, or events).
#include <mutex>

#include "agent.cuh"

namespace ois {
namespace {

constexpr int kMaxRows = OIS_STATIC_MLP_MAX_ROWS;
constexpr int kH0 = 512, kH1 = 512, kH2 = 128;                 // widths of the models on this path
constexpr size_t kScratchFloats = (size_t)kMaxRows * (kH0 + kH1 + kH2);

// In host code this name is a host-side stand-in, not a device address: only cudaGetSymbolAddress below yields the
// pointer a kernel may use.
static __device__ __align__(256) float g_mlp_scratch[kScratchFloats];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
float* g_scratch_ptr[kMaxDevices];   // per device, resolved once with that device current

// Device address of g_mlp_scratch on the CURRENT device (each device holds its own copy of the array).
int scratch_for_current_device(float** out) {
  int dev = -1;
  OIS_CHECK_CUDA(cudaGetDevice(&dev));
  OIS_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_ptr[dev]) {
    float* p = nullptr;
    OIS_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_mlp_scratch));
    OIS_CHECK_ARG(p != nullptr, "cudaGetSymbolAddress returned a null device address");
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return OIS_OK;
}

}  // namespace
}  // namespace ois

using namespace ois;

extern "C" int ois_mlp_forward_static(const OisMlpShape* shape, const float* params, const float* x, float* y,
                                      int64_t n_rows, void* stream) {
  OIS_CHECK_ARG(shape && params && x && y, "null argument");
  OIS_CHECK_ARG(n_rows >= 0, "negative n_rows");
  if (n_rows > kMaxRows) {
    set_error("ois_mlp_forward_static: n_rows %lld exceeds OIS_STATIC_MLP_MAX_ROWS = %d (the static scratch array is fixed; "
              "use ois_mlp_forward with an agent handle for more rows)", (long long)n_rows, kMaxRows);
    return OIS_ERR_INVALID_ARG;
  }
  OIS_CHECK_ARG(shape->in_dim > 0 && shape->out_dim > 0 && shape->h0 > 0 && shape->h1 > 0 && shape->h2 > 0, "bad shape");
  OIS_CHECK_ARG(shape->h0 <= kH0 && shape->h1 <= kH1 && shape->h2 <= kH2,
                "hidden sizes exceed the static scratch (512, 512, 128)");
  if (n_rows == 0) return OIS_OK;
  float* scratch = nullptr;
  if (int e = scratch_for_current_device(&scratch)) return e;
  float* h0 = scratch;
  float* h1 = h0 + (size_t)kMaxRows * kH0;
  float* h2 = h1 + (size_t)kMaxRows * kH1;
  return mlp_forward_fp32(*shape, params, x, shape->in_dim, h0, h1, h2, y, n_rows, (cudaStream_t)stream);
}
