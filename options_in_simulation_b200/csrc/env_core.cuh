// Device helpers of the fused environment step shared by env_step.cu (the stand-alone kernel) and mlp_tct.cu (the
// policy step as ONE kernel: the CTA that produced a tile's actions steps those envs): the reference's op-by-op
// arithmetic helpers, the constant-divisor division, the Philox draws, the skid-steer sub-steps and
// reset_idx (row A7r).  Moved here unchanged from env_step.cu.
#pragma once
#include "ois_common.cuh"

namespace ois {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float dvd(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float clipf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// ---------------------------------------------------------------------------------------------
// Correctly rounded division by a divisor that is constant for the launch (dt, action_scale, max_sim_steps).
// __fdiv_rn costs 11 instructions plus a branch per quotient (MUFU.RCP, FCHK, five FFMA) and falls into a subroutine
// for zero numerators; the kernel is issue-bound, and the step has ten such quotients per env.  With rc = RN(1/d):
//     q0 = RN(x rc);  r = x - d q0 (exact, one FMA);  q = RN(q0 + r rc)
// is RN(x / d) (Markstein 1990; Brisebarre, Muller & Raina 2004) for every x unless d is one of a few exceptional
// divisors — so the HOST checks each divisor once, exhaustively over all 2^23 significands of x (the result is
// scale-invariant), before the kernel is allowed to use it (const_div_exact below); otherwise `exact` is 0 and the
// kernel divides with __fdiv_rn.  Zero numerators return x (IEEE: +-0 / d = +-0 for d > 0).  Not covered: infinite,
// NaN or subnormal x — none of which a finite rollout state produces.
// ---------------------------------------------------------------------------------------------
struct ConstDiv {
  float d, rc;
  int exact;
};
__device__ __forceinline__ float cdiv(float x, const ConstDiv& c) {
  if (!c.exact) return __fdiv_rn(x, c.d);
  const float q0 = __fmul_rn(x, c.rc);
  const float r = __fmaf_rn(-c.d, q0, x);
  const float q = __fmaf_rn(r, c.rc, q0);
  return x == 0.0f ? x : q;
}
struct EnvDerived {
  ConstDiv dt, action_scale, max_sim_steps;
};

// Philox2x32-10 (Salmon et al. 2011, Random123): the per-step Bernoulli switch needs ONE uniform per env, and the
// four-lane generator costs 85 instructions for it; this one costs 25.  counter = (env, step offset), key = folded seed.
__device__ __forceinline__ uint32_t philox2x32(uint32_t c0, uint32_t c1, uint32_t key) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p = (uint64_t)0xD256D193u * c0;
    c0 = (uint32_t)(p >> 32) ^ key ^ c1;
    c1 = (uint32_t)p;
    key += 0x9E3779B9u;
  }
  return c0;
}

// torch.clamp(1.0 - (sim_steps.float() / max_sim_steps), min=0.0)   KAMMA_medium.py:627-628; `decay` is that value
__device__ __forceinline__ float gamma_assist(int mode, float constant, float decay) {
  return mode == OIS_GAMMA_DECAY ? decay : constant;
}

struct Skid {
  float px, py, yaw, vx, vy, r;
  float w[4];
};

// ---------------------------------------------------------------------------------------------
// `decimation` linearly-implicit sub-steps of the planar skid-steer model under constant torques.
// Equations: oracle/dynamics_twin.py (increment form).  Wheel order fl, rl, fr, rr:
// x = (+a, -a, +a, -a), y = (+b, +b, -b, -b).
//
// The kernel is issue-bound, not HBM-bound (profiles/r01_ncu_full_summary.txt), so the sub-step is written
// for instruction count: the two wheels of a side travel as one packed float2 (lane x = front, y = rear)
// through Blackwell's FFMA2 / FMUL2 / FADD2; reciprocals are MUFU.RCP + one Newton step (2 FMA, no IEEE
// slow path); the slip norm uses MUFU.RSQ directly.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 bc(float s) { return make_float2(s, s); }
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
#ifdef OIS_DIAG_NOMUFU   // timing diagnostic only (wrong results): what the MUFU pipe costs the sub-steps
__device__ __forceinline__ float mufu_rcp(float x) { return fmaf(x, 0.001f, 0.5f); }
__device__ __forceinline__ float mufu_rsq(float x) { return fmaf(x, 0.001f, 0.25f); }
#else
__device__ __forceinline__ float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float mufu_rsq(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif
// 1/x for two lanes: the MUFU.RCP result (<= 1 ulp) as is; OIS_ENV_RCP_NEWTON adds one Newton step r <- r + r (1 - x r)
#ifndef OIS_ENV_RCP_NEWTON
#define OIS_ENV_RCP_NEWTON 0
#endif
__device__ __forceinline__ float2 rcp2(float2 x) {
  const float2 r = make_float2(mufu_rcp(x.x), mufu_rcp(x.y));
#if OIS_ENV_RCP_NEWTON
  return fma2(r, fma2(neg2(x), r, bc(1.0f)), r);
#else
  return r;
#endif
}
__device__ __forceinline__ float rcp1(float x) {
  const float r = mufu_rcp(x);
#if OIS_ENV_RCP_NEWTON
  return fmaf(r, fmaf(-x, r, 1.0f), r);
#else
  return r;
#endif
}

struct SideOut { float2 al, be, k, g, phi; };

// one SIDE of the vehicle (lateral position y): its front (x = +a) and rear (x = -a) wheels in the two lanes.  Pairing by
// side keeps every packed operand in the registers it already lives in — wheel speeds and torques are stored
// (fl, rl, fr, rr), so (fl, rl) and (fr, rr) are adjacent — where pairing by axle cost a dozen register moves per sub-step.
// The sub-steps' time follows their count of FMA-pipe instructions, not the MUFU count and not the issue slots
// (profiles/r02p_variant_probe.txt, r02p_pair_probe.txt), so products are shared where the algebra allows:
// g = (Iw/h + c) k / E is be (Iw/h + c) / R, and sum k_i l_i is taken per axle after the sides are added (skid_steps).
__device__ __forceinline__ SideOut side(float2 w, float2 tau, float nu /* -u of this side */, float2 l2e /* l^2 + eps^2 */,
                                        float2 R2, float2 RR2, float2 Iwh2, float2 IwhR2 /* (Iw/h + c) / R */, float2 nc2,
                                        float2 mu2) {
  SideOut o;
  const float2 sx = fma2(R2, w, bc(nu));                   // R w - u
  const float2 n2 = fma2(sx, sx, l2e);
  const float2 k = mul2(make_float2(mufu_rsq(n2.x), mufu_rsq(n2.y)), mu2);
  const float2 inv = rcp2(fma2(RR2, k, Iwh2));             // 1 / (Iw/h + c + R^2 k)
  const float2 Rk = mul2(R2, k);
  o.al = mul2(fma2(neg2(Rk), sx, fma2(nc2, w, tau)), inv);  // (tau - c w - R k sx) / E
  o.be = mul2(Rk, inv);
  o.g = mul2(o.be, IwhR2);
  o.phi = mul2(k, fma2(R2, o.al, sx));
  o.k = k;
  return o;
}

// What the sub-steps need that depends on neither the torques nor the loop: four IEEE divisions and sincos(yaw).  Split
// off so that a caller that has the state before it has the torques (the one-launch policy step, mlp_tct.cu) can evaluate
// it while it waits; the arithmetic is the same either way.
struct SkidPre { float m_h, J_h, Iw_h, IwR_h, sn, cs; };
__device__ __forceinline__ SkidPre skid_prepare(float yaw, const OisEnvParams& p) {
  SkidPre q;
  const float h = p.sim_dt;
  q.m_h = p.mass / h;
  q.J_h = p.yaw_inertia / h;
  q.Iw_h = p.wheel_inertia / h + p.wheel_damping;
  q.IwR_h = q.Iw_h / p.wheel_radius;
  sincosf(yaw, &q.sn, &q.cs);
  return q;
}

__device__ __forceinline__ void skid_steps(Skid& s, const float tau[4], const OisEnvParams& p, const SkidPre& pre) {
  const float h = p.sim_dt, R = p.wheel_radius, c = p.wheel_damping;
  const float a = p.half_wheelbase, b = p.half_track;
  const float m = p.mass;
  const float m_h = pre.m_h, J_h = pre.J_h, Iw_h = pre.Iw_h;
  const float eps2 = p.slip_eps * p.slip_eps, aa = a * a, bb = b * b;
  const float wmax = p.wheel_speed_max;
  const float2 R2 = bc(R), RR2 = bc(R * R), Iwh2 = bc(Iw_h), IwhR2 = bc(pre.IwR_h), nc2 = bc(-c), mu2 = bc(p.mu_n);
  float2 wL = make_float2(s.w[0], s.w[1]), wR = make_float2(s.w[2], s.w[3]);   // left / right side, (front, rear)
  const float2 tL = make_float2(tau[0], tau[1]), tR = make_float2(tau[2], tau[3]);
  float sn = pre.sn, cs = pre.cs;
  float sx_w = 0.f, sy_w = 0.f;
#pragma unroll 1
  for (int it = 0; it < p.decimation; ++it) {
    const float vx = s.vx, vy = s.vy, r = s.r;
    const float2 l = make_float2(fmaf(r, a, vy), fmaf(-r, a, vy));     // lateral contact speed: front, rear
    const float2 l2e = fma2(l, l, bc(eps2));
    const SideOut Lf = side(wL, tL, fmaf(r, b, -vx), l2e, R2, RR2, Iwh2, IwhR2, nc2, mu2);     // y = +b: u = vx - r b
    const SideOut Rt = side(wR, tR, fmaf(-r, b, -vx), l2e, R2, RR2, Iwh2, IwhR2, nc2, mu2);    // y = -b: u = vx + r b
    const float2 Sg = add2(Lf.g, Rt.g), Dg = add2(Lf.g, neg2(Rt.g)), Sk = add2(Lf.k, Rt.k);
    const float2 Sph = add2(Lf.phi, Rt.phi), Dph = add2(Lf.phi, neg2(Rt.phi)), Skl = mul2(Sk, l);   // per axle: (k_l + k_r) l
    const float gs = Sg.x + Sg.y, ks = Sk.x + Sk.y;
    const float A00 = m_h + gs, A11 = m_h + ks;
    const float A22 = fmaf(bb, gs, fmaf(aa, ks, J_h));
    const float A02 = -b * (Dg.x + Dg.y);                    // -sum y_i g_i
    const float A12 = a * (Sk.x - Sk.y);                     //  sum x_i k_i
    const float b0 = fmaf(m * r, vy, Sph.x + Sph.y);
    const float b1 = -fmaf(m * r, vx, Skl.x + Skl.y);
    const float b2 = -fmaf(a, Skl.x - Skl.y, b * (Dph.x + Dph.y));
    const float2 iA = rcp2(make_float2(A00, A11));
    const float e0 = A02 * iA.x, e1 = A12 * iA.y;
    const float s22 = fmaf(-A12, e1, fmaf(-A02, e0, A22));
    const float dr = fmaf(-e1, b1, fmaf(-e0, b0, b2)) * rcp1(s22);
    const float dvx = fmaf(-A02, dr, b0) * iA.x;
    const float dvy = fmaf(-A12, dr, b1) * iA.y;
    s.vx = vx + dvx;
    s.vy = vy + dvy;
    s.r = r + dr;
    // u' - u = dvx -+ dr b  (left, right)
    wL = add2(wL, fma2(Lf.be, bc(fmaf(-dr, b, dvx)), Lf.al));
    wR = add2(wR, fma2(Rt.be, bc(fmaf(dr, b, dvx)), Rt.al));
    wL = make_float2(clipf(wL.x, -wmax, wmax), clipf(wL.y, -wmax, wmax));
    wR = make_float2(clipf(wR.x, -wmax, wmax), clipf(wR.y, -wmax, wmax));
#ifdef OIS_DIAG_NOHEADING   // timing diagnostic only (wrong results)
    s.yaw += h * s.r;
    sx_w += s.vx;
    continue;
#endif
    // heading: rotate (cos, sin) by d = h*r' with a Taylor pair.  |d| <= 0.01 * R wmax / b = 0.3: the first dropped terms,
    // d^9 / 9! of the sine and d^8 / 8! of the cosine, are below 6e-11 and 2e-9 there (fp32 ulp of 1: 6e-8)
    const float d = h * s.r, d2 = d * d;
    const float sd = d * (1.f + d2 * (-1.f / 6.f + d2 * (1.f / 120.f + d2 * (-1.f / 5040.f))));
    const float cd = 1.f + d2 * (-0.5f + d2 * (1.f / 24.f + d2 * (-1.f / 720.f)));
    const float cn = cs * cd - sn * sd;
    sn = sn * cd + cs * sd;
    cs = cn;
    s.yaw += d;
    // world-frame displacement / h, summed over the sub-steps and scaled once after the loop
    sx_w = fmaf(s.vx, cs, fmaf(-s.vy, sn, sx_w));
    sy_w = fmaf(s.vx, sn, fmaf(s.vy, cs, sy_w));
  }
  s.px = fmaf(h, sx_w, s.px);
  s.py = fmaf(h, sy_w, s.py);
  s.w[0] = wL.x; s.w[1] = wL.y; s.w[2] = wR.x; s.w[3] = wR.y;
}
__device__ __forceinline__ void skid_steps(Skid& s, const float tau[4], const OisEnvParams& p) {
  skid_steps(s, tau, p, skid_prepare(s.yaw, p));
}

// warp-level sum
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// reset_idx draws in reference order: pos x, pos y, angle, cmd x, cmd y, heading, loop x  (KAMMA_medium.py:424-456,571)
struct ResetDraws {
  float u[8];   // [7] is a standard normal, used by OIS_CMD_BD only
};

__device__ __forceinline__ void apply_reset(const OisEnvParams& p, const ResetDraws& d, int64_t sim_after, Skid& s,
                                            float4& cmd) {
  // pose: base_init_state pos (0,0) + U(-0.5,0.5)^2 ; yaw = U(0, 2*pi)   KAMMA_medium.py:422-434
  s.px = add(0.0f, add(mul(1.0f, d.u[0]), -0.5f));
  s.py = add(0.0f, add(mul(1.0f, d.u[1]), -0.5f));
  s.yaw = add(mul(6.2831855f, d.u[2]), 0.0f);
  s.vx = 0.f; s.vy = 0.f; s.r = 0.f;                                     // :435-446
  s.w[0] = s.w[1] = s.w[2] = s.w[3] = 0.f;
  // commands :448-460
  const float cx = add(mul(p.x_range, d.u[3]), p.x_lo);
  const float cy = add(mul(p.y_range, d.u[4]), p.y_lo);
  const float ch = add(mul(p.h_range, d.u[5]), p.h_lo);
  const float keep = sqrtf(add(mul(cx, cx), mul(cy, cy))) > 0.25f ? 1.0f : 0.0f;
  cmd.y = mul(cy, keep);
  cmd.w = mul(ch, keep);
  // per-env command via sample_velocity_command :490-496
  float x;
  if (p.command_mode == OIS_CMD_RAND) {            // :569-575
    x = add(mul(p.x_range, d.u[6]), p.x_lo);
  } else if (p.command_mode == OIS_CMD_GV) {       // KAMMA_curr1.py:577-599, per-env progress
    const float progress = fminf(1.0f, dvd((float)sim_after, p.max_sim_steps));
    const float factor = add(0.1f, mul(0.9f, progress));
    const float c = (float)(((double)p.x_lo + (double)p.x_hi) / 2.0);
    const float xmin = sub(c, mul((float)((double)c - (double)p.x_lo), factor));
    const float xmax = add(c, mul((float)((double)p.x_hi - (double)c), factor));
    x = add(mul(sub(xmax, xmin), d.u[6]), xmin);
  } else if (p.command_mode == OIS_CMD_BD) {       // KAMMA_curr2.py:601-641 (bimodal "grid" sampling), per-env progress
    // two modes that drift from the centre of the range to its ends; rand < 0.5 picks the mode, torch.normal(mean, 0.05)
    // = randn * std + mean draws around it (the omega draws that follow are discarded by the reference: omega = 0)
    const float progress = fminf(1.0f, dvd((float)sim_after, p.max_sim_steps));
    const double c = ((double)p.x_lo + (double)p.x_hi) / 2.0;
    const float left = sub((float)c, mul(progress, (float)(c - (double)p.x_lo)));
    const float right = add((float)c, mul(progress, (float)((double)p.x_hi - c)));
    x = add(mul(d.u[7], 0.05f), d.u[6] < 0.5f ? left : right);
  } else {                                          // KAMMA_eval.py:545-550
    const float progress = clipf(dvd((float)sim_after, p.max_sim_steps), 0.0f, 1.0f);
    x = add(1.0f, mul(progress, 1.0f));
  }
  cmd.x = fmaxf(x, 0.0f);
  cmd.z = 0.0f;
}

__device__ __forceinline__ ResetDraws philox_reset_draws(uint64_t seed, uint64_t offset, int64_t env, uint4 first) {
  ResetDraws d;
  const uint4 second = Philox(seed)((uint64_t)env, offset * 4 + 1);
  d.u[0] = u01(first.y); d.u[1] = u01(first.z); d.u[2] = u01(first.w);
  d.u[3] = u01(second.x); d.u[4] = u01(second.y); d.u[5] = u01(second.z); d.u[6] = u01(second.w);
  const uint4 third = Philox(seed)((uint64_t)env, offset * 4 + 2);   // resets are rare: drawn whether or not BD needs it
  d.u[7] = normal2(third.x, third.y).x;
  return d;
}

// Per-step scalars: by value for a plain launch, from the device-resident OisStepCtl under CUDA-graph replay.
__device__ __forceinline__ uint64_t env_philox_offset(const OisEnvIO& io) {
  return io.ctl ? io.ctl->env_philox_offset : io.philox_offset;
}
__device__ __forceinline__ int env_flag_slot(const OisEnvIO& io) { return io.ctl ? io.ctl->flag_slot : io.flag_slot; }
// first row of this step's ring slot, in rows of the flattened (memory_size * N, dim) view
__device__ __forceinline__ int64_t replay_row_base(const OisEnvIO& io, int64_t n) {
  return io.ctl ? io.ctl->replay_row * n : 0;
}

// launch-constant divisors of a step, checked on the host (env_step.cu: const_div_exact)
EnvDerived make_env_derived(const OisEnvParams* params);

}  // namespace ois
