// Shared device/host helpers for libois_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "ois_b200.h"

namespace ois {

void set_error(const char* fmt, ...);

#define OIS_CHECK_ARG(cond, msg)                         \
  do {                                                   \
    if (!(cond)) {                                       \
      ::ois::set_error("%s: %s", __func__, msg);         \
      return OIS_ERR_INVALID_ARG;                        \
    }                                                    \
  } while (0)

#define OIS_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      ::ois::set_error("%s: %s -> %s", __func__, #expr, cudaGetErrorString(_e));          \
      return OIS_ERR_CUDA;                                                                \
    }                                                                                     \
  } while (0)

#define OIS_CHECK_LAUNCH() OIS_CHECK_CUDA(cudaGetLastError())

inline int num_sms() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
}

// ---------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011), counter-based: throughput-mode RNG for the Bernoulli switch,
// resets, OU noise and replay sampling.  Parity tests bypass it by supplying the draws.
// ---------------------------------------------------------------------------------------------
struct Philox {
  uint32_t k0, k1;
  __host__ __device__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    uint64_t p = (uint64_t)a * b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
  }
  __host__ __device__ inline uint4 operator()(uint64_t ctr_lo, uint64_t ctr_hi) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32), c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      uint32_t h0, l0, h1, l1;
      mulhilo(0xD2511F53u, c0, h0, l0);
      mulhilo(0xCD9E8D57u, c2, h1, l1);
      c0 = h1 ^ c1 ^ a;
      c1 = l1;
      c2 = h0 ^ c3 ^ b;
      c3 = l0;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

// [0,1) with 24 random bits, the same mapping torch.rand uses for float32
__host__ __device__ inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// Box-Muller pair from two 32-bit words
__device__ inline float2 normal2(uint32_t a, uint32_t b) {
  float u1 = ((float)(a >> 8) + 1.0f) * (1.0f / 16777216.0f);  // (0,1]
  float u2 = u01(b);
  float r = sqrtf(-2.0f * logf(u1));
  float s, c;
  sincospif(2.0f * u2, &s, &c);
  return make_float2(r * c, r * s);
}

// ---------------------------------------------------------------------------------------------
// Programmatic dependent launch (sm_90+): a kernel launched with launch_pdl() may start while the previous kernel in
// the stream is still running; it must execute pdl_wait() before touching anything an earlier kernel wrote, and the
// earlier kernel lets it go with pdl_trigger().  Rule used throughout: every thread does pdl_wait() and only then
// pdl_trigger(), so whatever a kernel reads before ITS wait was written no later than by its grand-parent, which had
// completed before the parent triggered.  A kernel that never triggers behaves as a normal stream predecessor.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// diagnostics (profiles/update_phases.py): remaining launches launch_pdl may still issue; < 0 = unlimited.  With a
// budget of k only the first k launches of a sequence run, so timing k = 1, 2, ... gives every launch's share of a chain.
int& launch_budget();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  int& budget = launch_budget();
  if (budget == 0) return cudaSuccess;
  if (budget > 0) --budget;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  static const bool no_pdl = getenv("OIS_NO_PDL") != nullptr;   // diagnostics: plain stream order
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// The one place the per-step scalars of a graph-replayed rollout step are derived (see OisStepCtl in ois_b200.h); run by a
// single thread.  The noise scale follows ddpg.py:247-251 in double, as the host path (agent.py: DDPG.act) evaluates it
// before narrowing to float.
// The block's contents for the NEXT policy step.  Split from the store so that a kernel can evaluate it (an fp64 division,
// a 64-bit modulo, a dozen dependent loads: ~1.6 us for one thread) while it waits for something else, and its last CTA
// only has to store the result.
struct StepCtlNext {
  int64_t step;
  uint64_t env_philox_offset, act_philox_offset;
  int64_t replay_row;
  double timestep;
  int32_t flag_slot;
  float noise_scale;
};
__device__ inline StepCtlNext step_ctl_next(const OisStepCtl* c) {
  StepCtlNext nx;
  nx.step = c->step + 1;
  nx.env_philox_offset = c->env_philox_offset + 1;
  nx.act_philox_offset = c->act_philox_offset + 1;
  nx.replay_row = c->replay_rows > 0 ? (c->replay_row + 1) % c->replay_rows : 0;
  nx.flag_slot = (c->flag_slot + 1) % 3;
  const double t = c->timestep + 1.0;
  nx.timestep = t;
  nx.noise_scale = t <= c->expl_timesteps
      ? (float)((1.0 - t / c->expl_timesteps) * (c->expl_initial_scale - c->expl_final_scale) + c->expl_final_scale)
      : -1.0f;
  return nx;
}
__device__ inline void step_ctl_store(OisStepCtl* c, const StepCtlNext& nx) {
  c->step = nx.step;
  c->env_philox_offset = nx.env_philox_offset;
  c->act_philox_offset = nx.act_philox_offset;
  c->replay_row = nx.replay_row;
  c->flag_slot = nx.flag_slot;
  c->timestep = nx.timestep;
  c->noise_scale = nx.noise_scale;
}
__device__ inline void step_ctl_advance(OisStepCtl* c) { step_ctl_store(c, step_ctl_next(c)); }

// streaming 16-byte accesses: every state field is touched once per step
__device__ inline float4 ld_stream4(const float* p) { return __ldcs(reinterpret_cast<const float4*>(p)); }
__device__ inline void st_stream4(float* p, const float4& v) { __stcs(reinterpret_cast<float4*>(p), v); }

}  // namespace ois
