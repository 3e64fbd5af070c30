// OrnsteinUhlenbeckNoise.sample + ddpg.py:247-255 (scale, add, clamp) for ONE action element, shared by the
// standalone explore kernel (fp32 path) and the epilogue of the tcgen05 actor kernel, so both paths draw
// identical noise.  `i` = row * act_dim + column.
#pragma once
#include "ois_common.cuh"

namespace ois {

// Per-step scalars: by value for a plain launch, from the device-resident OisStepCtl under CUDA-graph replay.
__device__ __forceinline__ uint64_t act_philox_offset(const OisActIO& io) {
  return io.ctl ? io.ctl->act_philox_offset : io.philox_offset;
}
__device__ __forceinline__ float act_noise_scale(const OisActIO& io) { return io.ctl ? io.ctl->noise_scale : io.noise_scale; }

__device__ __forceinline__ float explore_elem(const OisActIO& io, int64_t i, float action, float& nz, bool& have) {
  float z;
  if (io.normals) {
    z = io.normals[i];
  } else {
    const uint4 r = Philox(io.philox_seed)((uint64_t)(i >> 2), act_philox_offset(io) * 2);
    const float2 n01 = normal2(r.x, r.y), n23 = normal2(r.z, r.w);
    const int k = (int)(i & 3);
    z = k == 0 ? n01.x : k == 1 ? n01.y : k == 2 ? n23.x : n23.y;
  }
  const float st = io.ou_state[i];
  const float ns = __fadd_rn(st, __fadd_rn(__fmul_rn(-st, io.theta), __fmul_rn(io.sigma, z)));   // state += -state*theta + sigma*N
  io.ou_state[i] = ns;
  if (act_noise_scale(io) >= 0.f) {
    nz = __fmul_rn(__fmul_rn(io.base_scale, ns), act_noise_scale(io));
    have = true;
    return fminf(fmaxf(__fadd_rn(action, nz), io.clip_low), io.clip_high);
  }
  return action;
}

// The same for one whole row when act_dim == 4 (one Philox call serves the row: counter = row, lanes 0..3 —
// exactly the draws explore_elem makes for i = 4*row + k).  Accumulates the applied noises into (mx, mn, sm).
__device__ __forceinline__ void explore_row4(const OisActIO& io, int64_t row, float (&a)[4], float& mx, float& mn, float& sm) {
  float z[4];
  if (io.normals) {
    const float4 n4 = *reinterpret_cast<const float4*>(io.normals + 4 * row);
    z[0] = n4.x; z[1] = n4.y; z[2] = n4.z; z[3] = n4.w;
  } else {
    const uint4 r = Philox(io.philox_seed)((uint64_t)row, act_philox_offset(io) * 2);
    const float2 n01 = normal2(r.x, r.y), n23 = normal2(r.z, r.w);
    z[0] = n01.x; z[1] = n01.y; z[2] = n23.x; z[3] = n23.y;
  }
  float4 st4 = *reinterpret_cast<const float4*>(io.ou_state + 4 * row);
  float st[4] = {st4.x, st4.y, st4.z, st4.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    st[k] = __fadd_rn(st[k], __fadd_rn(__fmul_rn(-st[k], io.theta), __fmul_rn(io.sigma, z[k])));
    if (act_noise_scale(io) >= 0.f) {
      const float nz = __fmul_rn(__fmul_rn(io.base_scale, st[k]), act_noise_scale(io));
      mx = fmaxf(mx, nz); mn = fminf(mn, nz); sm += nz;
      a[k] = fminf(fmaxf(__fadd_rn(a[k], nz), io.clip_low), io.clip_high);
    }
  }
  *reinterpret_cast<float4*>(io.ou_state + 4 * row) = make_float4(st[0], st[1], st[2], st[3]);
}

// OU update of one whole row for act_dim in {1, 2, 4}, WITHOUT touching the action: advances io.ou_state and returns the
// scaled noises nz[k] to add (applied == false: timestep past the exploration horizon, nothing is added).  Bit-identical
// to explore_elem for i = act_dim*row + k (the row's elements share one Philox counter block because act_dim divides 4).
// Lets the tensor-core actor prepare the noise while the MLP is still in flight.
__device__ __forceinline__ void explore_row_prepare(const OisActIO& io, int64_t row, int A, float (&nz)[4], bool& applied) {
  const int64_t i0 = row * A;
  float z[4] = {0.f, 0.f, 0.f, 0.f};
  if (io.normals) {
#pragma unroll
    for (int k = 0; k < 4; ++k) if (k < A) z[k] = io.normals[i0 + k];
  } else {
    const uint4 r = Philox(io.philox_seed)((uint64_t)(i0 >> 2), act_philox_offset(io) * 2);
    const int l0 = (int)(i0 & 3);
    if (A == 4) {
      const float2 n01 = normal2(r.x, r.y), n23 = normal2(r.z, r.w);
      z[0] = n01.x; z[1] = n01.y; z[2] = n23.x; z[3] = n23.y;
    } else {   // one Box-Muller pair covers the row: lanes (l0, l0+1) for A == 2, lane l0 for A == 1
      const float2 nn = (l0 & 2) ? normal2(r.z, r.w) : normal2(r.x, r.y);
      if (A == 2) { z[0] = nn.x; z[1] = nn.y; }
      else z[0] = (l0 & 1) ? nn.y : nn.x;
    }
  }
  applied = act_noise_scale(io) >= 0.f;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    nz[k] = 0.f;
    if (k < A) {
      const float st = io.ou_state[i0 + k];
      const float ns = __fadd_rn(st, __fadd_rn(__fmul_rn(-st, io.theta), __fmul_rn(io.sigma, z[k])));
      io.ou_state[i0 + k] = ns;
      if (applied) nz[k] = __fmul_rn(__fmul_rn(io.base_scale, ns), act_noise_scale(io));
    }
  }
}

__device__ __forceinline__ void atomic_max_f(float* addr, float v) {
  if (v >= 0.f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMin(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_min_f(float* addr, float v) {
  if (v >= 0.f) atomicMin(reinterpret_cast<int*>(addr), __float_as_int(v));
  else atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// warp-reduce (max, min, sum) of the applied noises and fold into io.noise_stats (ddpg.py:258-260)
__device__ __forceinline__ void noise_stats_warp(const OisActIO& io, float mx, float mn, float sm) {
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    sm += __shfl_xor_sync(0xffffffffu, sm, o);
  }
  if ((threadIdx.x & 31) == 0 && mx != -INFINITY) {
    atomic_max_f(io.noise_stats + 0, mx);
    atomic_min_f(io.noise_stats + 1, mn);
    atomicAdd(io.noise_stats + 2, sm);
  }
}

}  // namespace ois
