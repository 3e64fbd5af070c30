// DDPG agent entry points: act = scaler -> actor MLP -> OU noise -> clamp;
// update = scaler(train) x2 -> TD target -> critic step -> policy step -> Adam x2 -> polyak x2.
// Reference: src/my_agents/ddpg.py:215-268 (act), :357-444 (_update); models
// train/train_kamma_4d_nofifo_random_RLIL.py:21-52; skrl pieces restated in oracle/ref_harness/skrl_shim.py.
#include <new>
#include "agent.cuh"
#include "mlp_fp32.cuh"
#include "explore.cuh"
#include "update_tc.cuh"
#include "comm.cuh"

namespace ois {

// ------------------------------------------------------------------------------------------- layout
MlpOffsets mlp_offsets(const OisMlpShape& s) {
  MlpOffsets o;
  const int dims[5] = {s.in_dim, s.h0, s.h1, s.h2, s.out_dim};
  int64_t off = 0;
  for (int l = 0; l < 4; ++l) {
    o.w[l] = off;
    off += (int64_t)dims[l + 1] * dims[l];
    o.b[l] = off;
    off += dims[l + 1];
    o.in[l] = dims[l];
    o.out[l] = dims[l + 1];
  }
  o.total = off;
  return o;
}

// ------------------------------------------------------------------------------------------- kernels
// RunningStandardScaler._compute without train: clamp((x - mean.float()) / (sqrt(var.float()) + eps), +-5)
__global__ void scaler_apply_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n, int dim,
                                    const double* __restrict__ mean, const double* __restrict__ var, int ldy) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * dim) return;
  const int64_t row = i / dim;
  const int c = (int)(i - row * dim);
  const float mu = (float)mean[c], sd = sqrtf((float)var[c]);
  const float v = __fdiv_rn(__fsub_rn(x[i], mu), __fadd_rn(sd, 1e-8f));
  y[row * ldy + c] = fminf(fmaxf(v, -5.0f), 5.0f);
}

// skrl RunningStandardScaler._parallel_variance for one column: merge a batch (mean, unbiased variance, B rows) into the
// running (mu, vr, cur).  Explicitly rounded operations (no compiler-chosen FMA contraction), so that every caller —
// scaler_train_block and the fused batch preparation of the tensor-core update — produces the same bits.
__device__ __forceinline__ void scaler_merge(double& mu, double& vr, double& cur, double in_mean, double in_var, int B) {
  const double n = (double)B;
  const double tot = __dadd_rn(cur, n);
  const double delta = __dsub_rn(in_mean, mu);
  const double cross = __ddiv_rn(__dmul_rn(__dmul_rn(__dmul_rn(delta, delta), cur), n), tot);
  const double M2 = __dadd_rn(__dadd_rn(__dmul_rn(vr, cur), __dmul_rn(in_var, n)), cross);
  mu = __dadd_rn(mu, __ddiv_rn(__dmul_rn(delta, n), tot));
  vr = __ddiv_rn(M2, tot);
  cur = tot;
}

// RunningStandardScaler(x, train=True) on a [B, dim] batch, one CTA of 256 threads: batch mean / unbiased variance
// (torch.mean / torch.var are float32 results; we accumulate in fp64 and round once), parallel-variance merge into
// the fp64 running statistics, then standardise with the UPDATED statistics.  All columns are swept together (one
// global-memory round trip per pass instead of one per column); the fp64 summation order per column is fixed:
// thread-strided partial sums -> xor-shuffle tree -> the eight warp totals in warp order.
__device__ void scaler_train_block(const float* __restrict__ x, float* __restrict__ y, int B, int dim,
                                   double* __restrict__ mean, double* __restrict__ var, double* __restrict__ count, int ldy) {
  __shared__ double s_sum[8][16];
  __shared__ float s_mu[16], s_sd[16];
  __shared__ double s_bmean[16];
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  double a[16];
  // pass 1: mean
#pragma unroll
  for (int c = 0; c < 16; ++c) a[c] = 0.0;
  for (int r = t; r < B; r += 256) {
#pragma unroll
    for (int c = 0; c < 16; ++c)
      if (c < dim) a[c] += (double)x[(size_t)r * dim + c];
  }
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    if (c < dim) {
      for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
      if (lane == 0) s_sum[wp][c] = a[c];
    }
  }
  __syncthreads();
  if (t < dim) {
    double m = 0.0;
    for (int w = 0; w < 8; ++w) m += s_sum[w][t];
    s_bmean[t] = m / (double)B;
  }
  __syncthreads();
  // pass 2: centred sum of squares
#pragma unroll
  for (int c = 0; c < 16; ++c) a[c] = 0.0;
  for (int r = t; r < B; r += 256) {
#pragma unroll
    for (int c = 0; c < 16; ++c)
      if (c < dim) { const double d = (double)x[(size_t)r * dim + c] - s_bmean[c]; a[c] = __fma_rn(d, d, a[c]); }
  }
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    if (c < dim) {
      for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
      if (lane == 0) s_sum[wp][c] = a[c];
    }
  }
  __syncthreads();
  if (t < dim) {
    double m2 = 0.0;
    for (int w = 0; w < 8; ++w) m2 += s_sum[w][t];
    const double in_mean = (double)(float)s_bmean[t];                      // torch.mean -> float32
    const double in_var = (double)(float)(m2 / (double)(B - 1));           // torch.var (unbiased) -> float32
    double nm = mean[t], nv = var[t], cur = count[0];
    scaler_merge(nm, nv, cur, in_mean, in_var, B);
    mean[t] = nm;
    var[t] = nv;
    s_mu[t] = (float)nm;
    s_sd[t] = sqrtf((float)nv);
  }
  __syncthreads();
  if (t == 0) count[0] += (double)B;   // every reader of count[0] is past the barrier above
  for (int i = t; i < B * dim; i += 256) {
    const int r = i / dim, c = i - r * dim;
    const float v = __fdiv_rn(__fsub_rn(x[i], s_mu[c]), __fadd_rn(s_sd[c], 1e-8f));
    y[(size_t)r * ldy + c] = fminf(fmaxf(v, -5.0f), 5.0f);
  }
  __syncthreads();                      // y, the statistics and count are complete for whoever runs next in this CTA
}

__global__ void __launch_bounds__(256)
scaler_train_kernel(const float* __restrict__ x, float* __restrict__ y, int B, int dim, double* __restrict__ mean,
                    double* __restrict__ var, double* __restrict__ count, int ldy) {
  scaler_train_block(x, y, B, dim, mean, var, count, ldy);
}

// copy [B, w] block into columns [col0, col0+w) of a wider row-major matrix
__global__ void copy_cols_kernel(const float* __restrict__ src, int lds, float* __restrict__ dst, int ldd, int col0,
                                 int B, int w) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * w) return;
  const int r = i / w, c = i - r * w;
  dst[(size_t)r * ldd + col0 + c] = src[(size_t)r * lds + c];
}

// target_values = rewards + discount * not(terminated | truncated) * target_q      ddpg.py:391-396
__global__ void td_target_kernel(const float* __restrict__ r, const uint8_t* __restrict__ te,
                                 const uint8_t* __restrict__ tr, const float* __restrict__ q2, float discount,
                                 float* __restrict__ y, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const float nd = (te[i] | tr[i]) ? 0.0f : 1.0f;
  y[i] = __fadd_rn(r[i], __fmul_rn(__fmul_rn(discount, nd), q2[i]));
}

// critic loss = mean((q - y)^2), dq = 2 (q - y) / B; also target stats.   one CTA
// With r != null the TD target (ddpg.py:391-396) is computed here first; with dz != null the gradient is also written as
// column 0 of a tile-blocked bf16 [B, 16] matrix (the tensor-core backward's input, update_tc.cu).
__global__ void __launch_bounds__(256)
critic_loss_kernel(const float* __restrict__ q, float* __restrict__ y, float* __restrict__ dq, int B,
                   float* __restrict__ stats, const float* __restrict__ r, const uint8_t* __restrict__ te,
                   const uint8_t* __restrict__ tr, const float* __restrict__ q2, float discount, uint8_t* __restrict__ dz) {
  __shared__ float s_sum[8], s_ysum[8], s_ymax[8], s_ymin[8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  float a = 0.f, ys = 0.f, ymx = -INFINITY, ymn = INFINITY;
  for (int i = t; i < B; i += 256) {
    float yi;
    if (r) {
      const float nd = (te[i] | tr[i]) ? 0.0f : 1.0f;
      yi = __fadd_rn(r[i], __fmul_rn(__fmul_rn(discount, nd), q2[i]));
      y[i] = yi;
    } else {
      yi = y[i];
    }
    const float d = q[i] - yi;
    a += d * d;
    const float g = 2.0f * d / (float)B;
    dq[i] = g;
    if (dz) {
      *reinterpret_cast<uint4*>(dz + tcu_chunk_off(i, 0, 2, 128)) = make_uint4((uint32_t)tcu_bf16_bits(g), 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(dz + tcu_chunk_off(i, 1, 2, 128)) = make_uint4(0u, 0u, 0u, 0u);
    }
    ys += yi;
    ymx = fmaxf(ymx, yi);
    ymn = fminf(ymn, yi);
  }
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    ys += __shfl_xor_sync(0xffffffffu, ys, o);
    ymx = fmaxf(ymx, __shfl_xor_sync(0xffffffffu, ymx, o));
    ymn = fminf(ymn, __shfl_xor_sync(0xffffffffu, ymn, o));
  }
  if (lane == 0) { s_sum[wp] = a; s_ysum[wp] = ys; s_ymax[wp] = ymx; s_ymin[wp] = ymn; }
  __syncthreads();
  if (t == 0 && stats) {
    float A = 0.f, Y = 0.f, MX = -INFINITY, MN = INFINITY;
    for (int w = 0; w < 8; ++w) { A += s_sum[w]; Y += s_ysum[w]; MX = fmaxf(MX, s_ymax[w]); MN = fminf(MN, s_ymin[w]); }
    stats[0] = A / (float)B;
    stats[5] = MX; stats[6] = MN; stats[7] = Y / (float)B;
  }
}

__device__ __forceinline__ float bf16_lo_f(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi_f(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
__device__ __forceinline__ uint32_t pack_bf16_pair(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
// eight columns of d2 = acc * elu'(h): h = the saved bf16 layer-3 activations (elu' = 1 for h > 0, else h + 1)
__device__ __forceinline__ uint4 d2_chunk(const float (&acc)[8], const uint4 h) {
  const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
  uint32_t o[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float ha = bf16_lo_f(hw[j]), hb = bf16_hi_f(hw[j]);
    o[j] = pack_bf16_pair(acc[2 * j] * (ha > 0.f ? 1.f : ha + 1.f), acc[2 * j + 1] * (hb > 0.f ? 1.f : hb + 1.f));
  }
  return make_uint4(o[0], o[1], o[2], o[3]);
}

// Tensor-core update, ddpg.py:391-407 in ONE launch of B / 16 CTAs: TD target, MSE loss, d(loss)/dq as the bf16 dz column
// AND the first data gradient d2 = (dz W4) * elu'(h2) — with one output the K = 16 GEMM that produced d2 degenerates to
// bf16(g_r) * bf16(W4[0][n]), an exact fp32 product, so this is bit-identical to that launch and removes it from the
// dependent chain.  Thread (row, 8-column chunk); CTA 0 also reduces the loss / target statistics over the whole batch.
__global__ void __launch_bounds__(256)
critic_loss_tc_kernel(const float* __restrict__ q, float* __restrict__ y, float* __restrict__ dq, int B, float* __restrict__ stats,
                      const float* __restrict__ r, const uint8_t* __restrict__ te, const uint8_t* __restrict__ tr,
                      const float* __restrict__ q2, float discount, uint8_t* __restrict__ dz, uint8_t* __restrict__ d2,
                      const uint8_t* __restrict__ h2, const uint8_t* __restrict__ w4) {
  __shared__ float s_sum[8], s_ysum[8], s_ymax[8], s_ymin[8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  {
    const int row = blockIdx.x * 16 + (t >> 4), c8 = t & 15;
    if (row < B) {
      const float nd = (te[row] | tr[row]) ? 0.0f : 1.0f;
      const float yi = __fadd_rn(r[row], __fmul_rn(__fmul_rn(discount, nd), q2[row]));
      const float g = 2.0f * (q[row] - yi) / (float)B;
      const uint32_t gb = (uint32_t)tcu_bf16_bits(g);
      if (c8 < 2) *reinterpret_cast<uint4*>(dz + tcu_chunk_off(row, c8, 2, 128)) = make_uint4(c8 == 0 ? gb : 0u, 0u, 0u, 0u);
      const float gf = __uint_as_float(gb << 16);
      const uint4 w = *reinterpret_cast<const uint4*>(w4 + c8 * 256);                    // row 0 of W4, columns 8 c8 ..
      const uint4 h = *reinterpret_cast<const uint4*>(h2 + tcu_chunk_off(row, c8, 16, 128));
      const float acc[8] = {gf * bf16_lo_f(w.x), gf * bf16_hi_f(w.x), gf * bf16_lo_f(w.y), gf * bf16_hi_f(w.y),
                            gf * bf16_lo_f(w.z), gf * bf16_hi_f(w.z), gf * bf16_lo_f(w.w), gf * bf16_hi_f(w.w)};
      *reinterpret_cast<uint4*>(d2 + tcu_chunk_off(row, c8, 16, 128)) = d2_chunk(acc, h);
    }
  }
  if (blockIdx.x != 0) return;
  float a = 0.f, ys = 0.f, ymx = -INFINITY, ymn = INFINITY;
  for (int i = t; i < B; i += 256) {      // the same loop, order and arithmetic as critic_loss_kernel
    const float nd = (te[i] | tr[i]) ? 0.0f : 1.0f;
    const float yi = __fadd_rn(r[i], __fmul_rn(__fmul_rn(discount, nd), q2[i]));
    y[i] = yi;
    const float d = q[i] - yi;
    a += d * d;
    dq[i] = 2.0f * d / (float)B;
    ys += yi;
    ymx = fmaxf(ymx, yi);
    ymn = fminf(ymn, yi);
  }
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    ys += __shfl_xor_sync(0xffffffffu, ys, o);
    ymx = fmaxf(ymx, __shfl_xor_sync(0xffffffffu, ymx, o));
    ymn = fminf(ymn, __shfl_xor_sync(0xffffffffu, ymn, o));
  }
  if (lane == 0) { s_sum[wp] = a; s_ysum[wp] = ys; s_ymax[wp] = ymx; s_ymin[wp] = ymn; }
  __syncthreads();
  if (t == 0 && stats) {
    float A = 0.f, Y = 0.f, MX = -INFINITY, MN = INFINITY;
    for (int w = 0; w < 8; ++w) { A += s_sum[w]; Y += s_ysum[w]; MX = fmaxf(MX, s_ymax[w]); MN = fminf(MN, s_ymin[w]); }
    stats[0] = A / (float)B;
    stats[5] = MX; stats[6] = MN; stats[7] = Y / (float)B;
  }
}

// Tensor-core update, the actor's side of ddpg.py:423-429 in ONE launch of B / 16 CTAs: dz = d(loss)/d(action) * act'(a)
// (sigmoid' = a (1 - a), tanh' = 1 - a^2) as the bf16 [B, 16] backward input, the first data gradient
// d2 = (dz W4) * elu'(h2) (out_dim <= 8 terms per element, accumulated in fp32 in column order) and, in CTA 0, the
// policy-loss statistics of ddpg.py:423 (loss = -mean q, max / min / mean q) in policy_loss_kernel's summation order.
__global__ void __launch_bounds__(256)
actor_dz_kernel(const float* __restrict__ dx, int lddx, int col0, const float* __restrict__ act_out, int A, int act, int B,
                uint8_t* __restrict__ dz, uint8_t* __restrict__ d2, const uint8_t* __restrict__ h2,
                const uint8_t* __restrict__ w4, const float* __restrict__ q, float* __restrict__ stats) {
  __shared__ float s_sum[8], s_max[8], s_min[8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  {
    const int row = blockIdx.x * 16 + (t >> 4), c8 = t & 15;
    if (row < B) {
      float dzf[8];
      uint32_t dzb[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float x = 0.f;
        if (j < A) {
          x = dx[(size_t)row * lddx + col0 + j];
          const float av = act_out[(size_t)row * A + j];
          x = act == OIS_ACT_SIGMOID ? x * av * (1.f - av) : act == OIS_ACT_TANH ? x * (1.f - av * av) : x;
        }
        dzb[j] = (uint32_t)tcu_bf16_bits(x);
        dzf[j] = __uint_as_float(dzb[j] << 16);
      }
      if (c8 == 0)
        *reinterpret_cast<uint4*>(dz + tcu_chunk_off(row, 0, 2, 128)) =
            make_uint4(dzb[0] | (dzb[1] << 16), dzb[2] | (dzb[3] << 16), dzb[4] | (dzb[5] << 16), dzb[6] | (dzb[7] << 16));
      if (c8 == 1) *reinterpret_cast<uint4*>(dz + tcu_chunk_off(row, 1, 2, 128)) = make_uint4(0u, 0u, 0u, 0u);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j < A) {
          const uint4 w = *reinterpret_cast<const uint4*>(w4 + c8 * 256 + j * 16);       // row j of W4, columns 8 c8 ..
          acc[0] = fmaf(dzf[j], bf16_lo_f(w.x), acc[0]); acc[1] = fmaf(dzf[j], bf16_hi_f(w.x), acc[1]);
          acc[2] = fmaf(dzf[j], bf16_lo_f(w.y), acc[2]); acc[3] = fmaf(dzf[j], bf16_hi_f(w.y), acc[3]);
          acc[4] = fmaf(dzf[j], bf16_lo_f(w.z), acc[4]); acc[5] = fmaf(dzf[j], bf16_hi_f(w.z), acc[5]);
          acc[6] = fmaf(dzf[j], bf16_lo_f(w.w), acc[6]); acc[7] = fmaf(dzf[j], bf16_hi_f(w.w), acc[7]);
        }
      }
      const uint4 h = *reinterpret_cast<const uint4*>(h2 + tcu_chunk_off(row, c8, 16, 128));
      *reinterpret_cast<uint4*>(d2 + tcu_chunk_off(row, c8, 16, 128)) = d2_chunk(acc, h);
    }
  }
  if (blockIdx.x != 0 || q == nullptr) return;
  float a = 0.f, mx = -INFINITY, mn = INFINITY;
  for (int k = t; k < B; k += 256) {
    a += q[k];
    mx = fmaxf(mx, q[k]);
    mn = fminf(mn, q[k]);
  }
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if (lane == 0) { s_sum[wp] = a; s_max[wp] = mx; s_min[wp] = mn; }
  __syncthreads();
  if (t == 0 && stats) {
    float S = 0.f, MX = -INFINITY, MN = INFINITY;
    for (int w = 0; w < 8; ++w) { S += s_sum[w]; MX = fmaxf(MX, s_max[w]); MN = fminf(MN, s_min[w]); }
    stats[1] = -S / (float)B;
    stats[2] = MX; stats[3] = MN; stats[4] = S / (float)B;
  }
}

// policy loss = -mean(q_pi), dq = -1/B; also Q stats.   one CTA
__global__ void __launch_bounds__(256)
policy_loss_kernel(const float* __restrict__ q, float* __restrict__ dq, int B, float* __restrict__ stats,
                   uint8_t* __restrict__ dz) {
  __shared__ float s_sum[8], s_max[8], s_min[8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, lane = t & 31, wp = t >> 5;
  float a = 0.f, mx = -INFINITY, mn = INFINITY;
  for (int i = t; i < B; i += 256) {
    a += q[i];
    mx = fmaxf(mx, q[i]);
    mn = fminf(mn, q[i]);
    dq[i] = -1.0f / (float)B;
    if (dz) {
      *reinterpret_cast<uint4*>(dz + tcu_chunk_off(i, 0, 2, 128)) = make_uint4((uint32_t)tcu_bf16_bits(-1.0f / (float)B), 0u, 0u, 0u);
      *reinterpret_cast<uint4*>(dz + tcu_chunk_off(i, 1, 2, 128)) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if (lane == 0) { s_sum[wp] = a; s_max[wp] = mx; s_min[wp] = mn; }
  __syncthreads();
  if (t == 0 && stats) {
    float A = 0.f, MX = -INFINITY, MN = INFINITY;
    for (int w = 0; w < 8; ++w) { A += s_sum[w]; MX = fmaxf(MX, s_max[w]); MN = fminf(MN, s_min[w]); }
    stats[1] = -A / (float)B;
    stats[2] = MX; stats[3] = MN; stats[4] = A / (float)B;
  }
}

// d(pre-activation) of the actor's output layer from d(action): sigmoid' = a (1 - a), tanh' = 1 - a^2
__global__ void out_act_grad_kernel(const float* __restrict__ da, int ldda, int col0, const float* __restrict__ a,
                                    float* __restrict__ dz, int B, int A, int act) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * A) return;
  const int r = i / A, c = i - r * A;
  const float g = da[(size_t)r * ldda + col0 + c], v = a[i];
  dz[i] = act == OIS_ACT_SIGMOID ? g * v * (1.f - v) : act == OIS_ACT_TANH ? g * (1.f - v * v) : g;
}

// step_size = lr / (1 - beta1 ** step), sqrt(1 - beta2 ** step) of torch.optim.Adam, in double.  beta ** step by repeated
// squaring (<= 2 log2(step) double multiplications, relative error ~1e-15): every CTA pays this latency before its
// elementwise work, and the library pow() of a double was ~2 us of the Adam kernel's ~6.
__device__ __forceinline__ void adam_bias_corrections(int64_t t, double lr, double beta1, double beta2, float& step_size,
                                                      float& bc2_sqrt) {
  double p1 = 1.0, p2 = 1.0, x1 = beta1, x2 = beta2;
  for (int64_t e = t; e > 0; e >>= 1) {
    if (e & 1) { p1 = __dmul_rn(p1, x1); p2 = __dmul_rn(p2, x2); }
    x1 = __dmul_rn(x1, x1);
    x2 = __dmul_rn(x2, x2);
  }
  step_size = (float)(lr / (1.0 - p1));
  bc2_sqrt = (float)sqrt(1.0 - p2);
}

// one element of the single-tensor Adam step; returns the new parameter
__device__ __forceinline__ float adam_element(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                                              float g_sum, int64_t i, float grad_scale, double beta1,
                                              double beta2, float eps, float step_size, float bc2_sqrt) {
  const float w1 = (float)(1.0 - beta1), b2 = (float)beta2, w2 = (float)(1.0 - beta2);
  const float gi = g_sum * grad_scale;
  const float mi = m[i] + w1 * (gi - m[i]);            // exp_avg.lerp_(grad, 1 - beta1)
  const float vi = v[i] * b2 + w2 * gi * gi;           // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  m[i] = mi;
  v[i] = vi;
  const float pn = p[i] - step_size * (mi / denom);    // param.addcdiv_(exp_avg, denom, value=-step_size)
  p[i] = pn;
  return pn;
}

// torch.optim.Adam (single-tensor formulas), flat buffers; g is scaled by grad_scale (1/world_size) first.
// The step count may live on the device (CUDA-graph replay): bias corrections are then computed here.
// comm.world > 1: this launch IS the gradient all-reduce of ddpg.py:409-410 — barrier with the peers, then every element
// is the rank-ordered sum of all ranks' gradients read in place over NVLink (comm.cuh).
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const float* __restrict__ g, int64_t n,
            float grad_scale, double lr, double beta1, double beta2, float eps, int64_t step_host,
            const int64_t* __restrict__ step_dev, uint8_t* __restrict__ img, int in_dim, int out_dim,
            const __grid_constant__ CommArgs comm) {
  __shared__ float s_step_size, s_bc2_sqrt;
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  if (threadIdx.x == 0) adam_bias_corrections(step_dev ? step_dev[0] : step_host, lr, beta1, beta2, s_step_size, s_bc2_sqrt);
  const bool dp = comm.world > 1;
  if (dp) comm_barrier(comm);
  else __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float gs = dp ? comm_reduced(comm, i) : g[i];
    const float pn = adam_element(p, m, v, gs, i, grad_scale, beta1, beta2, eps, s_step_size, s_bc2_sqrt);
    if (img) tcu_img_store(img, i, pn, in_dim, out_dim);  // keep the bf16 tensor-core image current (update_tc.cu)
  }
  if (dp) comm_finish(comm);
}

// both target networks in one launch, each optionally with its bf16 tensor-core image kept current
struct PolyakPair {
  float* pt[2];
  const float* p[2];
  int64_t n[2];
  uint8_t* img[2];
  int in_dim[2], out_dim[2];
};
// Phase 2 of the tensor-core update in ONE launch: ddpg.py:438 Adam step of the actor and ddpg.py:443-444 polyak of both
// target networks.  Element i < n[0]: the actor's Adam step, then its target's polyak with the parameter just computed
// (the same value the separate polyak launch would have read back); element n[0] + j: the critic target's polyak (the
// critic was stepped in phase 1).  Same formulas as adam_kernel / polyak_kernel; the bf16 images are kept current.
struct AdamPolyak {
  PolyakPair pp;            // pt / p / n / img / dims of (actor, critic); p[0] is the actor being stepped
  float *m, *v;
  const float* g;
  uint8_t* actor_img;
  float grad_scale, eps;
  double lr, beta1, beta2;
  int64_t step_host;
  const int64_t* step_dev;
  CommArgs comm;            // world > 1: the actor-gradient all-reduce of ddpg.py:431-432 happens here (comm.cuh)
};
__global__ void __launch_bounds__(256)
adam_polyak_kernel(const __grid_constant__ AdamPolyak a, float tau) {
  __shared__ float s_step_size, s_bc2_sqrt;
  // No pdl_trigger here: this is the last kernel of an update and the writer of the actor's parameters and bf16 image.
  // Its stream successor is often the next policy step's actor kernel (mlp_tct.cu), which starts streaming that image
  // before its own griddepcontrol.wait; without a trigger that kernel cannot start until this one has completed.
  pdl_wait();
  if (threadIdx.x == 0)
    adam_bias_corrections(a.step_dev ? a.step_dev[0] : a.step_host, a.lr, a.beta1, a.beta2, s_step_size, s_bc2_sqrt);
  const bool dp = a.comm.world > 1;
  if (dp) comm_barrier(a.comm);
  else __syncthreads();
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < a.pp.n[0]) {
    float* p = const_cast<float*>(a.pp.p[0]);
    const float gs = dp ? comm_reduced(a.comm, i) : a.g[i];
    const float pn = adam_element(p, a.m, a.v, gs, i, a.grad_scale, a.beta1, a.beta2, a.eps, s_step_size, s_bc2_sqrt);
    if (a.actor_img) tcu_img_store(a.actor_img, i, pn, a.pp.in_dim[0], a.pp.out_dim[0]);
    const float t = __fadd_rn(__fmul_rn(a.pp.pt[0][i], 1.0f - tau), __fmul_rn(tau, pn));
    a.pp.pt[0][i] = t;
    if (a.pp.img[0]) tcu_img_store(a.pp.img[0], i, t, a.pp.in_dim[0], a.pp.out_dim[0]);
  } else if (i - a.pp.n[0] < a.pp.n[1]) {
    i -= a.pp.n[0];
    const float t = __fadd_rn(__fmul_rn(a.pp.pt[1][i], 1.0f - tau), __fmul_rn(tau, a.pp.p[1][i]));
    a.pp.pt[1][i] = t;
    if (a.pp.img[1]) tcu_img_store(a.pp.img[1], i, t, a.pp.in_dim[1], a.pp.out_dim[1]);
  }
  if (dp) comm_finish(a.comm);
}

// the tensor-core update's batch preparation in ONE CTA of 512 threads: ddpg.py:381-382 (scaler, train=True, on states
// then on next_states), then the bf16 network inputs:
//   x_critic  = [scaled states, sampled actions]   (critic input; its first D columns are also the actor's input)
//   x_scratch = [scaled next_states, 0]            (target actor input; the target actions land in columns D.. later)
//   x_policy  = [scaled states, 0]                 (policy-loss critic input; the actor's actions land in columns D..)
// The two scaler updates depend on each other only through the parallel-variance MERGE into the running statistics;
// the batch mean and the batch variance of each tensor do not.  So threads 0-255 reduce `states` and threads 256-511
// reduce `next_states` at the same time — each half with exactly the summation order of scaler_train_block (rows
// strided by 256 per thread, xor-shuffle tree, the eight warp totals in warp order), which keeps the statistics
// bit-identical to the fp32 path's — one thread per column then does the two merges back to back, and both tensors
// are standardised together.  Half the dependent stages of two scaler_train_block calls (23 us -> see DESIGN.md 3.1).
__global__ void __launch_bounds__(512)
tcu_batch_prep_kernel(const float* __restrict__ states, const float* __restrict__ next_states,
                      const float* __restrict__ actions, float* __restrict__ sn, float* __restrict__ s2n, int B, int D, int A,
                      double* __restrict__ mean, double* __restrict__ var, double* __restrict__ count,
                      uint8_t* __restrict__ x_critic, uint8_t* __restrict__ x_scratch, uint8_t* __restrict__ x_policy) {
  __shared__ double s_sum[2][8][16];
  __shared__ double s_bmean[2][16];
  __shared__ float s_mu[2][16], s_sd[2][16];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, half = t >> 8, tt = t & 255, lane = t & 31, wp = tt >> 5;
  const float* __restrict__ x = half ? next_states : states;
  float* __restrict__ y = half ? s2n : sn;
  if (mean) {
    double a[16];
    // pass 1: batch mean
#pragma unroll
    for (int c = 0; c < 16; ++c) a[c] = 0.0;
    for (int r = tt; r < B; r += 256) {
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < D) a[c] += (double)x[(size_t)r * D + c];
    }
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      if (c < D) {
        for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
        if (lane == 0) s_sum[half][wp][c] = a[c];
      }
    }
    __syncthreads();
    if (tt < D) {
      double m = 0.0;
      for (int w = 0; w < 8; ++w) m += s_sum[half][w][tt];
      s_bmean[half][tt] = m / (double)B;
    }
    __syncthreads();
    // pass 2: centred sum of squares
#pragma unroll
    for (int c = 0; c < 16; ++c) a[c] = 0.0;
    for (int r = tt; r < B; r += 256) {
#pragma unroll
      for (int c = 0; c < 16; ++c)
        if (c < D) { const double d = (double)x[(size_t)r * D + c] - s_bmean[half][c]; a[c] = __fma_rn(d, d, a[c]); }
    }
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      if (c < D) {
        for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
        if (lane == 0) s_sum[half][wp][c] = a[c];
      }
    }
    __syncthreads();
    if (t < D) {   // the two merges of skrl's _parallel_variance, in the reference's order: states, then next_states
      double mu = mean[t], vr = var[t], cur = count[0];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        double m2 = 0.0;
        for (int w = 0; w < 8; ++w) m2 += s_sum[h][w][t];
        const double in_mean = (double)(float)s_bmean[h][t];                  // torch.mean -> float32
        const double in_var = (double)(float)(m2 / (double)(B - 1));           // torch.var (unbiased) -> float32
        scaler_merge(mu, vr, cur, in_mean, in_var, B);
        s_mu[h][t] = (float)mu;
        s_sd[h][t] = sqrtf((float)vr);
      }
      mean[t] = mu;
      var[t] = vr;
    }
    __syncthreads();
    if (t == 0) count[0] += 2.0 * (double)B;   // every reader of count[0] is past the barrier above
    for (int i = tt; i < B * D; i += 256) {
      const int r = i / D, c = i - r * D;
      const float v = __fdiv_rn(__fsub_rn(x[i], s_mu[half][c]), __fadd_rn(s_sd[half][c], 1e-8f));
      y[i] = fminf(fmaxf(v, -5.0f), 5.0f);
    }
  } else {
    for (int i = tt; i < B * D; i += 256) y[i] = x[i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < B * 2; i += 512) {
    const int r = i >> 1, c8 = i & 1;
    uint16_t vc[8], vs[8], vp[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = c8 * 8 + j;
      const float xc = c < D ? sn[(size_t)r * D + c] : c < D + A ? actions[(size_t)r * A + (c - D)] : 0.f;
      const float xs = c < D ? s2n[(size_t)r * D + c] : 0.f;
      vc[j] = tcu_bf16_bits(xc);
      vs[j] = tcu_bf16_bits(xs);
      vp[j] = c < D ? vc[j] : (uint16_t)0;
    }
    auto pk = [](const uint16_t* v) {
      return make_uint4((uint32_t)v[0] | ((uint32_t)v[1] << 16), (uint32_t)v[2] | ((uint32_t)v[3] << 16),
                        (uint32_t)v[4] | ((uint32_t)v[5] << 16), (uint32_t)v[6] | ((uint32_t)v[7] << 16));
    };
    *reinterpret_cast<uint4*>(x_critic + tcu_chunk_off(r, c8, 2, 128)) = pk(vc);
    *reinterpret_cast<uint4*>(x_scratch + tcu_chunk_off(r, c8, 2, 128)) = pk(vs);
    *reinterpret_cast<uint4*>(x_policy + tcu_chunk_off(r, c8, 2, 128)) = pk(vp);
  }
}

// The same batch preparation with every thread's rows held in registers (B <= 256 RPT rows, D <= 8, A <= 8): each row is
// read from global memory ONCE, the two reduction passes and the standardisation run from registers, and the thread packs
// the bf16 network inputs of its own rows — no second and third sweep over global memory, no store-then-reload of the
// scaled states.  Summation order, merge and rounding are exactly tcu_batch_prep_kernel's (rows strided by 256 per
// thread in increasing order, xor-shuffle tree, warp totals in warp order), so the statistics stay bit-identical.
template <int RPT>
__global__ void __launch_bounds__(512)
tcu_batch_prep_reg_kernel(const float* __restrict__ states, const float* __restrict__ next_states,
                          const float* __restrict__ actions, float* __restrict__ sn, float* __restrict__ s2n, int B, int D, int A,
                          double* __restrict__ mean, double* __restrict__ var, double* __restrict__ count,
                          uint8_t* __restrict__ x_critic, uint8_t* __restrict__ x_scratch, uint8_t* __restrict__ x_policy) {
  __shared__ double s_sum[2][8][8];
  __shared__ double s_bmean[2][8];
  __shared__ float s_mu[2][8], s_sd[2][8];
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int t = threadIdx.x, half = t >> 8, tt = t & 255, lane = t & 31, wp = tt >> 5;
  const float* __restrict__ x = half ? next_states : states;
  float* __restrict__ y = half ? s2n : sn;
  float v[RPT][8], act[RPT][8];
  // the running statistics are needed only after both reduction passes: requested now, their latency hides behind them
  double mu_run = 0.0, vr_run = 0.0, cur_run = 0.0;
  if (mean && t < D) { mu_run = mean[t]; vr_run = var[t]; cur_run = count[0]; }
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int r = tt + 256 * k;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      v[k][c] = (r < B && c < D) ? x[(size_t)r * D + c] : 0.f;
      act[k][c] = (half == 0 && r < B && c < A) ? actions[(size_t)r * A + c] : 0.f;
    }
  }
  if (mean) {
    double a[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) a[c] = 0.0;
#pragma unroll
    for (int k = 0; k < RPT; ++k)
      if (tt + 256 * k < B) {
#pragma unroll
        for (int c = 0; c < 8; ++c) if (c < D) a[c] += (double)v[k][c];
      }
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c < D) {
        for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
        if (lane == 0) s_sum[half][wp][c] = a[c];
      }
    }
    __syncthreads();
    if (tt < D) {
      double m = 0.0;
      for (int w = 0; w < 8; ++w) m += s_sum[half][w][tt];
      s_bmean[half][tt] = m / (double)B;
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < 8; ++c) a[c] = 0.0;
#pragma unroll
    for (int k = 0; k < RPT; ++k)
      if (tt + 256 * k < B) {
#pragma unroll
        for (int c = 0; c < 8; ++c)
          if (c < D) { const double d = (double)v[k][c] - s_bmean[half][c]; a[c] = __fma_rn(d, d, a[c]); }
      }
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      if (c < D) {
        for (int o = 16; o > 0; o >>= 1) a[c] += __shfl_xor_sync(0xffffffffu, a[c], o);
        if (lane == 0) s_sum[half][wp][c] = a[c];
      }
    }
    __syncthreads();
    if (t < D) {   // the two merges of skrl's _parallel_variance, in the reference's order: states, then next_states
      double mu = mu_run, vr = vr_run, cur = cur_run;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        double m2 = 0.0;
        for (int w = 0; w < 8; ++w) m2 += s_sum[h][w][t];
        const double in_mean = (double)(float)s_bmean[h][t];                  // torch.mean -> float32
        const double in_var = (double)(float)(m2 / (double)(B - 1));           // torch.var (unbiased) -> float32
        scaler_merge(mu, vr, cur, in_mean, in_var, B);
        s_mu[h][t] = (float)mu;
        s_sd[h][t] = sqrtf((float)vr);
      }
      mean[t] = mu;
      var[t] = vr;
    }
    __syncthreads();
    if (t == 0) count[0] += 2.0 * (double)B;   // every reader of count[0] is past the barrier above
#pragma unroll
    for (int k = 0; k < RPT; ++k)
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c < D) v[k][c] = fminf(fmaxf(__fdiv_rn(__fsub_rn(v[k][c], s_mu[half][c]), __fadd_rn(s_sd[half][c], 1e-8f)), -5.0f), 5.0f);
  }
#pragma unroll
  for (int k = 0; k < RPT; ++k) {
    const int r = tt + 256 * k;
    if (r >= B) continue;
#pragma unroll
    for (int c = 0; c < 8; ++c) if (c < D) y[(size_t)r * D + c] = v[k][c];
    // columns 0..15 of the row: scaled states, then (critic input only) the sampled actions, then zeros
    uint16_t row[16], rowp[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      float xs = 0.f, xa = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j == c) xs = c < D ? v[k][j] : xs;
        if (j == c - D) xa = (c >= D && c < D + A) ? act[k][j] : xa;
      }
      rowp[c] = c < D ? tcu_bf16_bits(xs) : (uint16_t)0;
      row[c] = c < D ? rowp[c] : c < D + A ? tcu_bf16_bits(xa) : (uint16_t)0;
    }
    auto pk = [](const uint16_t* q) {
      return make_uint4((uint32_t)q[0] | ((uint32_t)q[1] << 16), (uint32_t)q[2] | ((uint32_t)q[3] << 16),
                        (uint32_t)q[4] | ((uint32_t)q[5] << 16), (uint32_t)q[6] | ((uint32_t)q[7] << 16));
    };
#pragma unroll
    for (int c8 = 0; c8 < 2; ++c8) {
      if (half == 0) {
        *reinterpret_cast<uint4*>(x_critic + tcu_chunk_off(r, c8, 2, 128)) = pk(row + 8 * c8);
        *reinterpret_cast<uint4*>(x_policy + tcu_chunk_off(r, c8, 2, 128)) = pk(rowp + 8 * c8);
      } else {
        *reinterpret_cast<uint4*>(x_scratch + tcu_chunk_off(r, c8, 2, 128)) = pk(rowp + 8 * c8);
      }
    }
  }
}

// Model.update_parameters: p_t.mul_(1 - polyak); p_t.add_(polyak * p)
__global__ void polyak_kernel(float* __restrict__ pt, const float* __restrict__ p, int64_t n, float tau) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  pt[i] = __fadd_rn(__fmul_rn(pt[i], 1.0f - tau), __fmul_rn(tau, p[i]));
}

// the random_timesteps branch (ddpg.py:229-230) or OU exploration on top of the actor output (explore.cuh)
__global__ void __launch_bounds__(256)
explore_kernel(const __grid_constant__ OisActIO io, int64_t n, int A) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float nz = 0.f;
  bool have = false;
  if (i < n * A) {
    if (io.random_phase) {   // Uniform(low, high).sample(): rand * (high - low) + low
      float u;
      if (io.uniforms) u = io.uniforms[i];
      else { const uint4 r = Philox(io.philox_seed)((uint64_t)(i >> 2), act_philox_offset(io) * 2 + 1); u = u01(((const uint32_t*)&r)[i & 3]); }
      io.actions[i] = __fadd_rn(__fmul_rn(u, __fsub_rn(io.clip_high, io.clip_low)), io.clip_low);
    } else if (io.ou_state) {
      io.actions[i] = explore_elem(io, i, io.actions[i], nz, have);
    }
  }
  if (io.noise_stats && !io.random_phase && io.ou_state && act_noise_scale(io) >= 0.f)
    noise_stats_warp(io, have ? nz : -INFINITY, have ? nz : INFINITY, have ? nz : 0.f);
}

// ------------------------------------------------------------------------------------------- host helpers
static inline int eg(int64_t n, int t) { return (int)((n + t - 1) / t); }

static int epi_for(int act) {
  return act == OIS_ACT_SIGMOID ? EPI_BIAS_SIGMOID : act == OIS_ACT_TANH ? EPI_BIAS_TANH : EPI_BIAS;
}

// y = mlp(x), saving hidden activations in h[0..2]
int mlp_forward_fp32(const OisMlpShape& sh, const float* P, const float* x, int ldx, float* h0, float* h1, float* h2,
                     float* y, int64_t rows, cudaStream_t s) {
  const MlpOffsets o = mlp_offsets(sh);
  const int M = (int)rows;
  if (int e = gemm_nt(x, P + o.w[0], h0, M, sh.h0, sh.in_dim, ldx, sh.in_dim, sh.h0, EPI_BIAS_ELU, P + o.b[0], nullptr, 0, s)) return e;
  if (int e = gemm_nt(h0, P + o.w[1], h1, M, sh.h1, sh.h0, sh.h0, sh.h0, sh.h1, EPI_BIAS_ELU, P + o.b[1], nullptr, 0, s)) return e;
  if (int e = gemm_nt(h1, P + o.w[2], h2, M, sh.h2, sh.h1, sh.h1, sh.h1, sh.h2, EPI_BIAS_ELU, P + o.b[2], nullptr, 0, s)) return e;
  return gemm_nt(h2, P + o.w[3], y, M, sh.out_dim, sh.h2, sh.h2, sh.h2, sh.out_dim, epi_for(sh.out_act), P + o.b[3], nullptr, 0, s);
}

// backward of the MLP given dz3 = d(loss)/d(pre-activation of the last layer) [B, out].
// G != null: parameter gradients are written (flat layout).  dx != null: d(loss)/d(input) [B, in].
static int mlp_backward_fp32(const OisMlpShape& sh, const float* P, float* G, const float* x, int ldx, const float* h0,
                             const float* h1, const float* h2, const float* dz3, float* d2, float* d1, float* d0,
                             float* dx, int B, cudaStream_t s) {
  const MlpOffsets o = mlp_offsets(sh);
  if (G) {
    if (int e = gemm_tn(dz3, h2, G + o.w[3], sh.out_dim, sh.h2, B, sh.out_dim, sh.h2, sh.h2, s)) return e;
    if (int e = col_sum(dz3, G + o.b[3], B, sh.out_dim, sh.out_dim, s)) return e;
  }
  if (int e = gemm_nn(dz3, P + o.w[3], d2, B, sh.h2, sh.out_dim, sh.out_dim, sh.h2, sh.h2, EPI_ELU_GRAD, nullptr, h2, sh.h2, s)) return e;
  if (G) {
    if (int e = gemm_tn(d2, h1, G + o.w[2], sh.h2, sh.h1, B, sh.h2, sh.h1, sh.h1, s)) return e;
    if (int e = col_sum(d2, G + o.b[2], B, sh.h2, sh.h2, s)) return e;
  }
  if (int e = gemm_nn(d2, P + o.w[2], d1, B, sh.h1, sh.h2, sh.h2, sh.h1, sh.h1, EPI_ELU_GRAD, nullptr, h1, sh.h1, s)) return e;
  if (G) {
    if (int e = gemm_tn(d1, h0, G + o.w[1], sh.h1, sh.h0, B, sh.h1, sh.h0, sh.h0, s)) return e;
    if (int e = col_sum(d1, G + o.b[1], B, sh.h1, sh.h1, s)) return e;
  }
  if (int e = gemm_nn(d1, P + o.w[1], d0, B, sh.h0, sh.h1, sh.h1, sh.h0, sh.h0, EPI_ELU_GRAD, nullptr, h0, sh.h0, s)) return e;
  if (G) {
    if (int e = gemm_tn(d0, x, G + o.w[0], sh.h0, sh.in_dim, B, sh.h0, ldx, sh.in_dim, s)) return e;
    if (int e = col_sum(d0, G + o.b[0], B, sh.h0, sh.h0, s)) return e;
  }
  if (dx) {
    if (int e = gemm_nn(d0, P + o.w[0], dx, B, sh.in_dim, sh.h0, sh.h0, sh.in_dim, sh.in_dim, EPI_NONE, nullptr, nullptr, 0, s)) return e;
  }
  return OIS_OK;
}

static int adam_step(float* p, float* m, float* v, const float* g, int64_t n, float grad_scale, double lr, double b1,
                     double b2, double eps, int64_t step, const int64_t* step_dev, uint8_t* img, const OisMlpShape& sh,
                     const CommArgs& comm, cudaStream_t s) {
  OIS_CHECK_CUDA(launch_pdl(adam_kernel, eg(n, 256), 256, 0, s, p, m, v, g, n, grad_scale, lr, b1, b2, (float)eps, step,
                            step_dev, img, (int)sh.in_dim, (int)sh.out_dim, comm));
  return OIS_OK;
}

// the exchange table of gradient `kind` when the caller set OisUpdateIO.comm for a world of more than one rank
static int comm_for(const OisUpdateIO* io, int kind, const float* grad, int64_t n, CommArgs& out) {
  out = CommArgs{};
  const OisComm* c = io->comm;
  if (!c || c->world <= 1) return OIS_OK;
  OIS_CHECK_ARG(c->connected, "OisUpdateIO.comm is not connected to its peers (ois_comm_connect)");
  OIS_CHECK_ARG(c->n[kind] == n, "OisUpdateIO.comm was created for other parameter counts");
  OIS_CHECK_ARG((const char*)grad == (const char*)c->block + c->off_grad[kind],
                "with OisUpdateIO.comm set the gradient buffers must be ois_comm_grad(comm, 0 / 1)");
  out = comm_args(c, kind);
  return OIS_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Deferred actor launch (rollout fusion).  Between ois_rollout_fuse_begin and ois_rollout_fuse_end (per host thread) a
// bf16 ois_agent_act that the fused kernel can serve is RECORDED instead of launched; the ois_env_step that follows
// and consumes its actions launches both as one kernel (rollout_try_fuse).  Anything else launches the recorded actor
// first, so stream order is what the caller wrote.
// ---------------------------------------------------------------------------------------------------------------
struct PendingAct {
  bool armed = false, have = false, has_scaler = false;
  OisAgent* a = nullptr;
  const float* params = nullptr;
  OisScaler scaler;
  OisActIO io;
  int64_t n = 0;
  cudaStream_t s = nullptr;
};
static thread_local PendingAct g_pend;

static int launch_pending() {
  if (!g_pend.have) return OIS_OK;
  g_pend.have = false;
  return tc_actor_forward(g_pend.a, g_pend.params, g_pend.has_scaler ? &g_pend.scaler : nullptr, &g_pend.io, g_pend.n, g_pend.s);
}

int rollout_try_fuse(const OisEnvParams* params, const OisEnvState* state, const OisEnvIO* io, int64_t n, cudaStream_t s) {
  if (!g_pend.have) return 0;
  OisAgent* a = g_pend.a;
  const bool fusable = g_pend.n == n && g_pend.s == s && io->actions == g_pend.io.actions && io->ep_stats == nullptr &&
                       io->comp0 == nullptr && !params->eval_sums && params->act_dim == a->actor.out_dim;
  if (!fusable) {
    const int e = launch_pending();
    return e ? e : 0;
  }
  g_pend.have = false;
  if (int e = tc_sync_image(a, g_pend.params, n, s)) return e;
  const int e = tct_rollout_forward(a, tcu_image_if_valid(a, TCU_ACTOR, g_pend.params), g_pend.params,
                                    g_pend.has_scaler ? &g_pend.scaler : nullptr, &g_pend.io, params, state, io, n, s);
  return e ? e : 1;
}

}  // namespace ois

using namespace ois;

extern "C" int ois_rollout_fuse_begin(void) {
  if (int e = launch_pending()) return e;
  g_pend.armed = true;
  return OIS_OK;
}

extern "C" int ois_rollout_fuse_end(void) {
  g_pend.armed = false;
  return launch_pending();
}

// ------------------------------------------------------------------------------------------- C ABI
extern "C" int ois_agent_create(OisAgent** out, const OisMlpShape* actor, const OisMlpShape* critic, int64_t max_rows,
                                int32_t batch_size, int32_t precision, void* stream) {
  OIS_CHECK_ARG(out && actor && critic, "null argument");
  OIS_CHECK_ARG(max_rows > 0 && batch_size > 0, "max_rows and batch_size must be positive");
  OIS_CHECK_ARG(precision == OIS_PRECISION_FP32 || precision == OIS_PRECISION_BF16, "unknown precision");
  OIS_CHECK_ARG(critic->in_dim == actor->in_dim + actor->out_dim, "critic input must be obs + act");
  OIS_CHECK_ARG(critic->out_dim == 1, "critic output must be 1");
  OIS_CHECK_ARG(actor->in_dim <= 16 && critic->in_dim <= 16 && actor->out_dim <= 8, "dims exceed the supported MLP family");
  OisAgent* a = new (std::nothrow) OisAgent();
  OIS_CHECK_ARG(a, "out of host memory");
  a->actor = *actor;
  a->critic = *critic;
  a->max_rows = max_rows;
  a->batch = batch_size;
  a->precision = precision;
  const int hmax0 = actor->h0 > critic->h0 ? actor->h0 : critic->h0;
  const int hmax1 = actor->h1 > critic->h1 ? actor->h1 : critic->h1;
  const int hmax2 = actor->h2 > critic->h2 ? actor->h2 : critic->h2;
  const int64_t B = batch_size;
  size_t bytes = 0;
  auto take = [&](int64_t n) { size_t off = bytes; bytes += ((size_t)n * sizeof(float) + 255) / 256 * 256; return off; };
  // update workspace (batch rows)
  const size_t o_sn = take(B * actor->in_dim), o_s2n = take(B * actor->in_dim), o_x = take(B * critic->in_dim);
  const size_t o_a = take(B * actor->out_dim), o_q = take(B), o_q2 = take(B), o_y = take(B), o_dq = take(B);
  const size_t o_ah0 = take(B * hmax0), o_ah1 = take(B * hmax1), o_ah2 = take(B * hmax2);
  const size_t o_ch0 = take(B * hmax0), o_ch1 = take(B * hmax1), o_ch2 = take(B * hmax2);
  const size_t o_d0 = take(B * hmax0), o_d1 = take(B * hmax1), o_d2 = take(B * hmax2);
  const size_t o_dx = take(B * critic->in_dim), o_dz = take(B * actor->out_dim);
  // act workspace (fp32 path): normalised states + three activation buffers for max_rows rows
  size_t o_xn = 0, o_h0 = 0, o_h1 = 0, o_h2 = 0;
  o_xn = take(max_rows * actor->in_dim);
  if (precision == OIS_PRECISION_FP32) {
    o_h0 = take(max_rows * actor->h0);
    o_h1 = take(max_rows * actor->h1);
    o_h2 = take(max_rows * actor->h2);
  }
  cudaError_t ce = cudaMalloc(&a->ws, bytes);
  if (ce != cudaSuccess) {
    delete a;
    set_error("ois_agent_create: cudaMalloc(%zu) -> %s", bytes, cudaGetErrorString(ce));
    return OIS_ERR_CUDA;
  }
  a->ws_bytes = bytes;
  char* base = (char*)a->ws;
  auto at = [&](size_t off) { return (float*)(base + off); };
  a->u_sn = at(o_sn); a->u_s2n = at(o_s2n); a->u_x = at(o_x); a->u_a = at(o_a); a->u_q = at(o_q); a->u_q2 = at(o_q2);
  a->u_y = at(o_y); a->u_dq = at(o_dq);
  a->ah0 = at(o_ah0); a->ah1 = at(o_ah1); a->ah2 = at(o_ah2);
  a->ch0 = at(o_ch0); a->ch1 = at(o_ch1); a->ch2 = at(o_ch2);
  a->d0 = at(o_d0); a->d1 = at(o_d1); a->d2 = at(o_d2); a->u_dx = at(o_dx); a->u_dz = at(o_dz);
  a->xn = at(o_xn);
  if (precision == OIS_PRECISION_FP32) { a->h0 = at(o_h0); a->h1 = at(o_h1); a->h2 = at(o_h2); }
  if (precision == OIS_PRECISION_BF16) {
    int e = tc_agent_init(a, (cudaStream_t)stream);
    if (!e) e = tcu_init(a);
    if (e) {
      tc_agent_free(a);
      cudaFree(a->ws);
      delete a;
      return e;
    }
  }
  *out = a;
  return OIS_OK;
}

extern "C" int ois_agent_destroy(OisAgent* a) {
  if (!a) return OIS_OK;
  tc_agent_free(a);
  tcu_free(a);
  if (a->ws) cudaFree(a->ws);
  delete a;
  return OIS_OK;
}

extern "C" int ois_agent_invalidate(OisAgent* a) {
  OIS_CHECK_ARG(a, "null agent");
  if (a->precision == OIS_PRECISION_BF16) { tc_invalidate_weights(a); tcu_invalidate(a, -1); }
  return OIS_OK;
}

extern "C" int ois_scaler_forward(const OisScaler* sc, const float* x, float* y, int64_t n_rows, int32_t dim,
                                  int32_t train, void* stream) {
  OIS_CHECK_ARG(sc && sc->running_mean && sc->running_variance && sc->current_count && x && y, "null argument");
  OIS_CHECK_ARG(dim > 0 && dim <= 16 && n_rows >= 0, "bad sizes");
  OIS_CHECK_ARG(!train || (n_rows >= 2 && n_rows <= (1 << 24)), "train needs 2 <= n_rows <= 2^24");
  if (n_rows == 0) return OIS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  if (train) scaler_train_kernel<<<1, 256, 0, s>>>(x, y, (int)n_rows, dim, sc->running_mean, sc->running_variance, sc->current_count, dim);
  else scaler_apply_kernel<<<eg(n_rows * dim, 256), 256, 0, s>>>(x, y, n_rows, dim, sc->running_mean, sc->running_variance, dim);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

extern "C" int ois_mlp_forward(OisAgent* a, const OisMlpShape* shape, const float* params, const float* x, float* y,
                               int64_t n_rows, void* stream) {
  OIS_CHECK_ARG(a && shape && params && x && y, "null argument");
  OIS_CHECK_ARG(n_rows >= 0 && n_rows <= a->max_rows, "n_rows exceeds the agent's max_rows");
  OIS_CHECK_ARG(a->precision == OIS_PRECISION_FP32, "ois_mlp_forward is the fp32 path; use ois_agent_act for bf16");
  OIS_CHECK_ARG(shape->h0 <= a->actor.h0 && shape->h1 <= a->actor.h1 && shape->h2 <= a->actor.h2, "hidden sizes exceed workspace");
  if (n_rows == 0) return OIS_OK;
  return mlp_forward_fp32(*shape, params, x, shape->in_dim, a->h0, a->h1, a->h2, y, n_rows, (cudaStream_t)stream);
}

extern "C" int ois_agent_sync_weights(OisAgent* a, const float* actor_params, int64_t n_rows, void* stream) {
  OIS_CHECK_ARG(a && actor_params, "null argument");
  if (a->precision != OIS_PRECISION_BF16) return OIS_OK;   // the fp32 path reads the master weights directly
  return tc_sync_image(a, actor_params, n_rows, (cudaStream_t)stream);
}

extern "C" int ois_agent_act(OisAgent* a, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                             int64_t n_rows, void* stream) {
  OIS_CHECK_ARG(a && actor_params && io, "null argument");
  OIS_CHECK_ARG(io->states && io->actions, "null states/actions");
  OIS_CHECK_ARG(n_rows >= 0 && n_rows <= a->max_rows, "n_rows exceeds the agent's max_rows");
  if (n_rows == 0) return OIS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  const int A = a->actor.out_dim, D = a->actor.in_dim;
  if (!io->random_phase) {
    if (a->precision == OIS_PRECISION_BF16) {
      if (g_pend.armed && !g_pend.have && tc_use_transposed(a, n_rows) && tct_can_fuse_env(n_rows) && A <= 4) {
        g_pend.have = true;               // recorded: the env step that consumes these actions launches both (rollout_try_fuse)
        g_pend.a = a;
        g_pend.params = actor_params;
        g_pend.has_scaler = scaler != nullptr;
        if (scaler) g_pend.scaler = *scaler;
        g_pend.io = *io;
        g_pend.n = n_rows;
        g_pend.s = s;
        return OIS_OK;
      }
      if (int e = launch_pending()) return e;
      return tc_actor_forward(a, actor_params, scaler, io, n_rows, s);   // exploration fused into its epilogue
    } else {
      const float* x = io->states;
      if (scaler && scaler->running_mean) {
        scaler_apply_kernel<<<eg(n_rows * D, 256), 256, 0, s>>>(io->states, a->xn, n_rows, D, scaler->running_mean,
                                                               scaler->running_variance, D);
        OIS_CHECK_LAUNCH();
        x = a->xn;
      }
      if (int e = mlp_forward_fp32(a->actor, actor_params, x, D, a->h0, a->h1, a->h2, io->actions, n_rows, s)) return e;
    }
  }
  if (io->random_phase || io->ou_state) {
    explore_kernel<<<eg(n_rows * A, 256), 256, 0, s>>>(*io, n_rows, A);
    OIS_CHECK_LAUNCH();
  }
  return OIS_OK;
}

extern "C" int ois_ddpg_update_phase(OisAgent* a, const OisScaler* scaler, const OisUpdateIO* io, int32_t phase,
                                     float grad_scale, void* stream) {
  OIS_CHECK_ARG(a && io, "null argument");
  OIS_CHECK_ARG(phase >= 0 && phase <= 2, "phase must be 0, 1 or 2");
  OIS_CHECK_ARG(io->actor && io->actor_target && io->critic && io->critic_target, "null parameter buffer");
  OIS_CHECK_ARG(io->actor_grad && io->critic_grad && io->actor_m && io->actor_v && io->critic_m && io->critic_v,
                "null optimizer buffer");
  cudaStream_t s = (cudaStream_t)stream;
  OIS_CHECK_ARG(io->batch_rows >= 0 && io->batch_rows <= a->batch, "batch_rows exceeds the handle's batch_size");
  const int B = io->batch_rows > 0 ? io->batch_rows : a->batch;
  const int D = a->actor.in_dim, A = a->actor.out_dim, CI = a->critic.in_dim;
  const MlpOffsets oa = mlp_offsets(a->actor), oc = mlp_offsets(a->critic);
  const bool has_scaler = scaler && scaler->running_mean;
  OIS_CHECK_ARG(!has_scaler || B >= 2, "the scaler's batch statistics need at least 2 rows");
  // OIS_PRECISION_BF16: every contraction below runs as a tcgen05 grouped GEMM (update_tc.cu) and the element-wise
  // pieces (fp64 scaler statistics, TD target, losses, Adam, polyak — the same arithmetic as the exact fp32 path) are
  // fused around them; every launch is a programmatic dependent launch, so prologues overlap the predecessor's tail
  const bool tc = a->precision == OIS_PRECISION_BF16;
  OIS_CHECK_ARG(!tc || a->tcu, "the bf16 tcgen05 update needs batch_size to be a multiple of 128 (use OIS_PRECISION_FP32 otherwise)");
  OIS_CHECK_ARG(!tc || B == a->batch, "the bf16 tcgen05 update runs full batches only (batch_rows must equal batch_size)");
  if (phase == 0) {
    OIS_CHECK_ARG(io->states && io->actions && io->rewards && io->next_states && io->terminated && io->truncated,
                  "null batch tensor");
    if (tc) {
      // ddpg.py:381-382 scaler(train=True) on states, then next_states; bf16 inputs of the critic and the target actor
      uint8_t* x_critic = tcu_x_buffer(a, TCU_ACTS_CRITIC);
      uint8_t* x_scratch = tcu_x_buffer(a, TCU_ACTS_SCRATCH);
      uint8_t* x_policy = tcu_x_buffer(a, TCU_ACTS_POLICY);
      double* const sm = has_scaler ? scaler->running_mean : (double*)nullptr;
      double* const sv = has_scaler ? scaler->running_variance : (double*)nullptr;
      double* const sc = has_scaler ? scaler->current_count : (double*)nullptr;
      static const bool prep_generic = getenv("OIS_PREP_GENERIC") != nullptr;      // diagnostics: the three-sweep kernel
      if (!prep_generic && D <= 8 && A <= 8 && B <= 512)        // every row of a thread in registers (agent.cu above)
        OIS_CHECK_CUDA(launch_pdl(tcu_batch_prep_reg_kernel<2>, 1, 512, 0, s, io->states, io->next_states, io->actions, a->u_sn,
                                  a->u_s2n, B, D, A, sm, sv, sc, x_critic, x_scratch, x_policy));
      else if (!prep_generic && D <= 8 && A <= 8 && B <= 1024)
        OIS_CHECK_CUDA(launch_pdl(tcu_batch_prep_reg_kernel<4>, 1, 512, 0, s, io->states, io->next_states, io->actions, a->u_sn,
                                  a->u_s2n, B, D, A, sm, sv, sc, x_critic, x_scratch, x_policy));
      else
        OIS_CHECK_CUDA(launch_pdl(tcu_batch_prep_kernel, 1, 512, 0, s, io->states, io->next_states, io->actions, a->u_sn, a->u_s2n,
                                  B, D, A, sm, sv, sc, x_critic, x_scratch, x_policy));
      // ddpg.py:385-386 target actor on next_states and ddpg.py:399 critic on (states, actions) are independent: they
      // advance layer by layer in the same grouped launches; the target actions land in the scratch input's columns D..
      {
        const int nets[2] = {TCU_ACTOR_T, TCU_CRITIC}, acts[2] = {TCU_ACTS_SCRATCH, TCU_ACTS_CRITIC};
        const float* P32[2] = {io->actor_target, io->critic};
        float* y[2] = {a->u_a, a->u_q};
        uint8_t* xo[2] = {x_scratch, nullptr};
        if (int e = tcu_forward(a, 2, nets, P32, acts, y, xo, D, s)) return e;
      }
      // ddpg.py:387-396 target critic on (next_states, target actions).  The actor's forward of ddpg.py:420 rides in the
      // same four launches: it reads only the scaled states and the actor's weights, neither of which the critic's
      // optimizer step (ddpg.py:416, phase 1) touches, so hoisting it above that step changes no value.  Its fp32
      // actions reuse u_a (the target actions there were consumed in bf16 form) and its bf16 actions land in x_policy
      {
        const int nets[2] = {TCU_CRITIC_T, TCU_ACTOR}, acts[2] = {TCU_ACTS_SCRATCH, TCU_ACTS_ACTOR};
        const float* P32[2] = {io->critic_target, io->actor};
        float* y[2] = {a->u_q2, a->u_a};
        uint8_t* xo[2] = {nullptr, x_policy};
        if (int e = tcu_forward(a, 2, nets, P32, acts, y, xo, D, s)) return e;
      }
      // TD target + MSE + d(loss)/dq as the backward's bf16 input (ddpg.py:391-407)
      // and the backward's first data gradient d2 (critic_loss_tc_kernel): the chain continues at layer 3
      OIS_CHECK_CUDA(launch_pdl(critic_loss_tc_kernel, (B + 15) / 16, 256, 0, s, (const float*)a->u_q, a->u_y, a->u_dq, B, io->stats,
                                io->rewards, io->terminated, io->truncated, (const float*)a->u_q2, io->discount,
                                tcu_x_buffer(a, TCU_DZ), tcu_d2_buffer(a), tcu_h2_buffer(a, TCU_ACTS_CRITIC),
                                tcu_w4_image(a, TCU_CRITIC)));
      return tcu_backward(a, TCU_CRITIC, io->critic_grad, TCU_ACTS_CRITIC, nullptr, false, true, s);
    }
    // ddpg.py:381-382  scaler(train=True) on states, then on next_states
    if (has_scaler) {
      scaler_train_kernel<<<1, 256, 0, s>>>(io->states, a->u_sn, B, D, scaler->running_mean, scaler->running_variance,
                                            scaler->current_count, D);
      scaler_train_kernel<<<1, 256, 0, s>>>(io->next_states, a->u_s2n, B, D, scaler->running_mean,
                                            scaler->running_variance, scaler->current_count, D);
      OIS_CHECK_LAUNCH();
    } else {
      OIS_CHECK_CUDA(cudaMemcpyAsync(a->u_sn, io->states, sizeof(float) * B * D, cudaMemcpyDeviceToDevice, s));
      OIS_CHECK_CUDA(cudaMemcpyAsync(a->u_s2n, io->next_states, sizeof(float) * B * D, cudaMemcpyDeviceToDevice, s));
    }
    // ddpg.py:385-396  target actor / target critic / TD target
    if (int e = mlp_forward_fp32(a->actor, io->actor_target, a->u_s2n, D, a->ah0, a->ah1, a->ah2, a->u_a, B, s)) return e;
    copy_cols_kernel<<<eg(B * D, 256), 256, 0, s>>>(a->u_s2n, D, a->u_x, CI, 0, B, D);
    copy_cols_kernel<<<eg(B * A, 256), 256, 0, s>>>(a->u_a, A, a->u_x, CI, D, B, A);
    if (int e = mlp_forward_fp32(a->critic, io->critic_target, a->u_x, CI, a->ch0, a->ch1, a->ch2, a->u_q2, B, s)) return e;
    td_target_kernel<<<eg(B, 256), 256, 0, s>>>(io->rewards, io->terminated, io->truncated, a->u_q2, io->discount, a->u_y, B);
    // ddpg.py:399-407  critic forward, MSE, backward
    copy_cols_kernel<<<eg(B * D, 256), 256, 0, s>>>(a->u_sn, D, a->u_x, CI, 0, B, D);
    copy_cols_kernel<<<eg(B * A, 256), 256, 0, s>>>(io->actions, A, a->u_x, CI, D, B, A);
    if (int e = mlp_forward_fp32(a->critic, io->critic, a->u_x, CI, a->ch0, a->ch1, a->ch2, a->u_q, B, s)) return e;
    critic_loss_kernel<<<1, 256, 0, s>>>(a->u_q, a->u_y, a->u_dq, B, io->stats, nullptr, nullptr, nullptr, nullptr, 0.f, nullptr);
    OIS_CHECK_LAUNCH();
    return mlp_backward_fp32(a->critic, io->critic, io->critic_grad, a->u_x, CI, a->ch0, a->ch1, a->ch2, a->u_dq, a->d2,
                             a->d1, a->d0, nullptr, B, s);
  }
  if (phase == 1) {
    // ddpg.py:416  critic optimizer step (bf16 path: the critic's tensor-core image is kept current by the same kernel)
    uint8_t* cimg = tc ? tcu_image_if_valid(a, TCU_CRITIC, io->critic) : nullptr;
    if (tc && !cimg) tcu_invalidate(a, TCU_CRITIC);
    CommArgs comm;
    if (int e = comm_for(io, 0, io->critic_grad, oc.total, comm)) return e;
    if (int e = adam_step(io->critic, io->critic_m, io->critic_v, io->critic_grad, oc.total, grad_scale, io->critic_lr,
                          io->beta1, io->beta2, io->adam_eps, io->adam_step, io->adam_step_dev, cimg, a->critic, comm, s)) return e;
    if (tc) {
      // ddpg.py:420-423 policy loss through the updated critic; the actor's forward already ran in phase 0 (its actions
      // are in x_policy's columns D.. and in u_a), so only the critic's forward on (states, actor actions) is left.
      // TCU_ACTS_POLICY shares the hidden-activation buffers of TCU_ACTS_CRITIC, which phase 0's backward is done with
      // Forward and data gradient through the critic as one chain (tcu_policy_pass): the loss gradient w.r.t. q is the
      // constant -1/B, so layer 3 of the forward already emits the first data gradient and layer 4 (q, read only by the
      // statistics) rides in the next launch.  The critic's parameter gradients of this pass are not computed: the
      // reference discards them at the next critic_optimizer.zero_grad() (ddpg.py:406).
      if (int e = tcu_policy_pass(a, TCU_CRITIC, io->critic, TCU_ACTS_POLICY, a->u_q, a->u_dx, s)) return e;
      // through the output activation into the actor: dz, the actor's first data gradient and the policy-loss statistics
      OIS_CHECK_ARG(A <= 8, "the fused actor output gradient handles out_dim <= 8");
      OIS_CHECK_CUDA(launch_pdl(actor_dz_kernel, (B + 15) / 16, 256, 0, s, (const float*)a->u_dx, CI, D, (const float*)a->u_a, A,
                                a->actor.out_act, B, tcu_x_buffer(a, TCU_DZ), tcu_d2_buffer(a),
                                tcu_h2_buffer(a, TCU_ACTS_ACTOR), tcu_w4_image(a, TCU_ACTOR), (const float*)a->u_q, io->stats));
      return tcu_backward(a, TCU_ACTOR, io->actor_grad, TCU_ACTS_ACTOR, nullptr, false, true, s);
    }
    // ddpg.py:420-429  policy loss through the (updated) critic
    if (int e = mlp_forward_fp32(a->actor, io->actor, a->u_sn, D, a->ah0, a->ah1, a->ah2, a->u_a, B, s)) return e;
    copy_cols_kernel<<<eg(B * D, 256), 256, 0, s>>>(a->u_sn, D, a->u_x, CI, 0, B, D);
    copy_cols_kernel<<<eg(B * A, 256), 256, 0, s>>>(a->u_a, A, a->u_x, CI, D, B, A);
    if (int e = mlp_forward_fp32(a->critic, io->critic, a->u_x, CI, a->ch0, a->ch1, a->ch2, a->u_q, B, s)) return e;
    policy_loss_kernel<<<1, 256, 0, s>>>(a->u_q, a->u_dq, B, io->stats, nullptr);
    OIS_CHECK_LAUNCH();
    // dgrad only through the critic (its parameter gradients from this pass are discarded by the reference at
    // the next critic_optimizer.zero_grad(), ddpg.py:406), then the actor's full backward
    if (int e = mlp_backward_fp32(a->critic, io->critic, nullptr, a->u_x, CI, a->ch0, a->ch1, a->ch2, a->u_dq, a->d2, a->d1,
                                  a->d0, a->u_dx, B, s)) return e;
    out_act_grad_kernel<<<eg(B * A, 256), 256, 0, s>>>(a->u_dx, CI, D, a->u_a, a->u_dz, B, A, a->actor.out_act);
    OIS_CHECK_LAUNCH();
    return mlp_backward_fp32(a->actor, io->actor, io->actor_grad, a->u_sn, D, a->ah0, a->ah1, a->ah2, a->u_dz, a->d2, a->d1,
                             a->d0, nullptr, B, s);
  }
  // phase 2: ddpg.py:438 actor optimizer step, :443-444 polyak
  uint8_t* aimg = tc ? tcu_image_if_valid(a, TCU_ACTOR, io->actor) : nullptr;
  CommArgs comm;
  if (int e = comm_for(io, 1, io->actor_grad, oa.total, comm)) return e;
  if (!tc)
    if (int e = adam_step(io->actor, io->actor_m, io->actor_v, io->actor_grad, oa.total, grad_scale, io->actor_lr, io->beta1,
                          io->beta2, io->adam_eps, io->adam_step, io->adam_step_dev, aimg, a->actor, comm, s)) return e;
  if (tc) {
    AdamPolyak ap;
    PolyakPair& pp = ap.pp;
    pp.pt[0] = io->actor_target; pp.p[0] = io->actor; pp.n[0] = oa.total;
    pp.pt[1] = io->critic_target; pp.p[1] = io->critic; pp.n[1] = oc.total;
    pp.img[0] = tcu_image_if_valid(a, TCU_ACTOR_T, io->actor_target);
    pp.img[1] = tcu_image_if_valid(a, TCU_CRITIC_T, io->critic_target);
    pp.in_dim[0] = a->actor.in_dim; pp.out_dim[0] = a->actor.out_dim;
    pp.in_dim[1] = a->critic.in_dim; pp.out_dim[1] = a->critic.out_dim;
    ap.m = io->actor_m; ap.v = io->actor_v; ap.g = io->actor_grad; ap.actor_img = aimg;
    ap.grad_scale = grad_scale; ap.eps = (float)io->adam_eps;
    ap.lr = io->actor_lr; ap.beta1 = io->beta1; ap.beta2 = io->beta2;
    ap.step_host = io->adam_step; ap.step_dev = io->adam_step_dev;
    ap.comm = comm;
    OIS_CHECK_CUDA(launch_pdl(adam_polyak_kernel, eg(oa.total + oc.total, 256), 256, 0, s, ap, io->polyak));
    tc_invalidate_weights(a);                       // the inference kernel's own stream image is rebuilt by the next act
    if (!aimg) tcu_invalidate(a, TCU_ACTOR);
    if (!pp.img[0]) tcu_invalidate(a, TCU_ACTOR_T);
    if (!pp.img[1]) tcu_invalidate(a, TCU_CRITIC_T);
    return OIS_OK;
  }
  polyak_kernel<<<eg(oa.total, 256), 256, 0, s>>>(io->actor_target, io->actor, oa.total, io->polyak);
  polyak_kernel<<<eg(oc.total, 256), 256, 0, s>>>(io->critic_target, io->critic, oc.total, io->polyak);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}
