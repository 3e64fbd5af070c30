// Fused environment step for TorqueDistributionTask.
//
// One thread per env; every state field is a [N,4] float array so each access is one coalesced 16-byte
// load/store (512 contiguous bytes per warp instruction); int64 counters are 8-byte accesses.  The ten
// physics sub-steps run in registers: state makes exactly one HBM round trip per policy step.
//
// Arithmetic contract
//   * the reference's tensor math (pre_physics_step / is_done / calculate_metrics / reset_idx /
//     get_observations) is evaluated op by op in the reference's order with non-contracting intrinsics
//     (__fmul_rn, __fadd_rn, __fdiv_rn ...), so it is bit-identical to CPU torch / the numpy oracle;
//   * the skid-steer sub-steps (no reference code exists: PhysX) use FMA contraction and MUFU
//     reciprocals; they are held to <= 1e-6 of the float64 twin (oracle/dynamics_twin.py).
#include <map>
#include <mutex>
#include <string.h>
#include <math.h>
#include "ois_common.cuh"
#include "env_core.cuh"
#include "agent.cuh"

namespace ois {

#ifndef OIS_ENV_THREADS
#define OIS_ENV_THREADS 128
#endif
constexpr int kEnvThreads = OIS_ENV_THREADS;
#ifndef OIS_ENV_MIN_BLOCKS
#define OIS_ENV_MIN_BLOCKS 5
#endif
constexpr int kEnvMinBlocks = OIS_ENV_MIN_BLOCKS;   // 5: <= 102 registers, 20 resident warps per SM (issue/latency-bound kernel)

// kStats / kEval are compile-time so that the common launch carries neither the 24 logging accumulators nor the
// eval-only state in registers.  Only what the sub-steps need is loaded before them; episode sums, the episode
// counter and the previous observation (replay append) are loaded after the physics.
// cp.async (LDGSTS) staging: global -> shared without passing through registers.  Each thread only ever reads the
// slots it filled itself, so cp.async.wait_group is the only synchronisation needed.
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory"); }

// Inputs of one env, staged per thread.  Group A = what the sub-steps depend on (double-buffered: the next
// grid-stride iteration's group A lands while this iteration computes); group B = what is only needed after the
// physics (issued a whole iteration ahead as well, single buffer, re-issued as soon as it has been consumed).
struct StageA {
  float4 pv[kEnvThreads], tp[kEnvThreads], wv[kEnvThreads], cmd[kEnvThreads], act[kEnvThreads];
  int64_t sim[kEnvThreads];
};
struct StageB {
  float4 sa[kEnvThreads], sb[kEnvThreads], se[kEnvThreads];
  int64_t eb[kEnvThreads];
  float2 obs[3][kEnvThreads];
};

// Diagnostics (not part of the public ABI; profiles/tct_trace.py): %globaltimer of the earliest CTA start [0], the earliest
// and latest return from griddepcontrol.wait [1], [2], and the latest CTA end [3] of the last launch.
// Compiled in only with -DOIS_ENV_TRACE (make EXTRA=-DOIS_ENV_TRACE): four atomics per CTA are visible at 6000 CTAs.
__device__ unsigned long long g_env_trace[4];
#ifndef OIS_ENV_TRACE
#define ENV_STAMP_MIN(k) do { } while (0)
#define ENV_STAMP_MAX(k) do { } while (0)
#else
#define ENV_STAMP_MIN(k) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); atomicMin(&g_env_trace[k], t_); } } while (0)
#define ENV_STAMP_MAX(k) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); atomicMax(&g_env_trace[k], t_); } } while (0)
#endif

template <bool kStats, bool kEval>
__global__ void __launch_bounds__(kEnvThreads, kEnvMinBlocks)
env_step_kernel(const __grid_constant__ OisEnvParams p, const __grid_constant__ OisEnvState st,
                const __grid_constant__ OisEnvIO io, const int64_t n, const __grid_constant__ EnvDerived dv) {
  __shared__ float s_stats[OIS_EP_STATS_LEN];
  // Programmatic dependent launch: this grid may have started while its stream predecessor (normally the actor kernel)
  // was still running.  The predecessor may be another env step (tests, dynamics-only loops), so nothing is read
  // before the wait; what overlaps is the launch itself.  The next kernel (the next actor step) is let go at once: it
  // reads nothing of ours before its own wait.
  // ... with one exception that is always safe: L2 prefetch hints for this thread's first env.  A hint moves no data
  // into the SM, so whatever the predecessor still writes is what the loads after the wait will see (L2 is the point of
  // coherence); it only makes those loads L2 hits when the state was last touched long ago (a cold cache, or an evicted
  // TLB entry, is ~2 us of latency in front of a ~5 us kernel).
  {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) {
      auto pf = [](const void* q) { asm volatile("prefetch.global.L2 [%0];" ::"l"(q)); };
      pf(st.pose_vx + 4 * j); pf(st.twist_prev + 4 * j); pf(st.wheel_speed + 4 * j); pf(st.commands + 4 * j);
      pf(st.sums_a + 4 * j); pf(st.sums_b + 4 * j); pf(st.sim_steps + j); pf(st.episode_buf + j);
      if (kEval) { pf(st.sums_eval + 4 * j); pf(st.old_torques + 4 * j); }
      if (io.replay_states != nullptr) pf(io.obs + 6 * j);
    }
  }
  ENV_STAMP_MIN(0);
  pdl_wait();
  ENV_STAMP_MIN(1);
  ENV_STAMP_MAX(2);
  pdl_trigger();
  if (kStats) {
    if (threadIdx.x < OIS_EP_STATS_LEN) s_stats[threadIdx.x] = 0.f;
    __syncthreads();
  }
  float acc[kStats ? OIS_EP_STATS_LEN : 1];
  if (kStats) {
#pragma unroll
    for (int k = 0; k < OIS_EP_STATS_LEN; ++k) acc[k] = 0.f;
  }
  int any_reset_local = 0;
  const bool zero_old = kEval && io.any_reset_flags != nullptr && io.any_reset_flags[env_flag_slot(io)] != 0;
  const bool replay = io.replay_states != nullptr;

  __shared__ __align__(16) StageA s_a[2];
  __shared__ __align__(16) StageB s_b;
  const int tid = threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  auto issue_a_state = [&](int buf, int64_t j) {
    if (j < n) {
      StageA& a = s_a[buf];
      cp_async16(&a.pv[tid], st.pose_vx + 4 * j);
      cp_async16(&a.tp[tid], st.twist_prev + 4 * j);
      cp_async16(&a.wv[tid], st.wheel_speed + 4 * j);
      cp_async16(&a.cmd[tid], st.commands + 4 * j);
      cp_async8(&a.sim[tid], st.sim_steps + j);
    }
  };
  auto issue_a_actions = [&](int buf, int64_t j) {
    if (j < n) {
      StageA& a = s_a[buf];
      if (p.act_dim == 4) cp_async16(&a.act[tid], io.actions + 4 * j);
      else if (p.act_dim == 2) cp_async8(&a.act[tid], io.actions + 2 * j);
      else cp_async4(&a.act[tid], io.actions + j);
    }
    cp_async_commit();
  };
  auto issue_a = [&](int buf, int64_t j) {
    issue_a_state(buf, j);
    issue_a_actions(buf, j);
  };
  auto issue_b = [&](int64_t j) {
    if (j < n) {
      cp_async16(&s_b.sa[tid], st.sums_a + 4 * j);
      cp_async16(&s_b.sb[tid], st.sums_b + 4 * j);
      cp_async8(&s_b.eb[tid], st.episode_buf + j);
      if (kEval) cp_async16(&s_b.se[tid], st.sums_eval + 4 * j);
      if (replay) {
        const float* o = io.obs + 6 * j;
        cp_async8(&s_b.obs[0][tid], o);
        cp_async8(&s_b.obs[1][tid], o + 2);
        cp_async8(&s_b.obs[2][tid], o + 4);
      }
    }
    cp_async_commit();
  };
  const int64_t i0 = (int64_t)blockIdx.x * blockDim.x + tid;
  issue_a(0, i0);
  issue_b(i0);
  int buf = 0;

  for (int64_t i = i0; i < n; i += stride, buf ^= 1) {
    // ------------------------------------------------------------------ group A: this env's now, the next one's in flight
    issue_a(buf ^ 1, i + stride);
    cp_async_wait<2>();               // pending at most: B(i), A(i + stride)
    const float4 pv = s_a[buf].pv[tid];
    const float4 tp = s_a[buf].tp[tid];
    const float4 wv = s_a[buf].wv[tid];
    float4 cmd = s_a[buf].cmd[tid];
    int64_t sim = s_a[buf].sim[tid];
    float a4[4];
    {
      const float4 a = s_a[buf].act[tid];
      if (p.act_dim == 4) { a4[0] = a.x; a4[1] = a.y; a4[2] = a.z; a4[3] = a.w; }            // KAMMA_medium.py:600
      else if (p.act_dim == 2) { a4[0] = a.x; a4[1] = a.x; a4[2] = a.y; a4[3] = a.y; }       // KA_DDPG_2D.py:624
      else { a4[0] = a4[1] = a4[2] = a4[3] = a.x; }                                          // KA_DDPG_1D.py:600
    }
    if (io.clip_actions > 0.f) {      // VecEnvRLGames.step clamps the actions before pre_physics_step (clipActions)
#pragma unroll
      for (int j = 0; j < 4; ++j) a4[j] = clipf(a4[j], -io.clip_actions, io.clip_actions);
    }
    const bool need_mask_draw = p.blend == OIS_BLEND_SWITCH && io.rand_mask == nullptr;

    // ------------------------------------------------------------------ pre_physics_step :604-636
    const float desired_v = cmd.x, desired_w = cmd.z;
    const float current_v = pv.w, current_w = tp.y;
    const float v_delta = sub(desired_v, current_v);
    const float omega_delta = sub(desired_w, current_w);
    const float lin = mul(p.kp, mul(p.vehicle_mass, cdiv(v_delta, dv.dt)));
    const float ang = mul(p.kp_omega, mul(p.vehicle_inertia, cdiv(omega_delta, dv.dt)));
    const float ac_left = sub(lin, ang), ac_right = add(lin, ang);
    const float crit[4] = {ac_left, ac_left, ac_right, ac_right};
    const float decay = fmaxf(sub(1.0f, cdiv((float)sim, dv.max_sim_steps)), 0.0f);
    const float g1 = gamma_assist(p.gamma1_mode, p.gamma1_const, decay);
    const float g2 = gamma_assist(p.gamma2_mode, p.gamma2_const, decay);
    float tq[4], wt[4];
    if (p.blend == OIS_BLEND_SWITCH) {          // :632-636
      const float rm = need_mask_draw ? u01(philox2x32((uint32_t)i, (uint32_t)env_philox_offset(io),
                                                       (uint32_t)io.philox_seed ^ (uint32_t)(io.philox_seed >> 32) ^ (uint32_t)(i >> 32)))
                                      : __ldcs(io.rand_mask + i);
      const float mask = rm < g1 ? 1.0f : 0.0f;
      const float nmask = sub(1.0f, mask);
#pragma unroll
      for (int j = 0; j < 4; ++j) tq[j] = add(mul(mask, crit[j]), mul(nmask, mul(a4[j], p.action_scale)));
    } else {                                    // KA_DDPG_1D.py:632-633
      const float ng = sub(1.0f, g1);
#pragma unroll
      for (int j = 0; j < 4; ++j) tq[j] = add(mul(mul(ng, a4[j]), p.action_scale), mul(g1, crit[j]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) wt[j] = clipf(tq[j], -p.torque_clip, p.torque_clip);   // :649
    float guiding;
    {
      float acc2 = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {             // KA_DDPG_1D.py:642 (pre-clip) | KAMMA_medium.py:656 (post-clip)
        const float d = p.guiding_post_clip ? sub(wt[j], crit[j]) : sub(mul(a4[j], p.action_scale), crit[j]);
        acc2 = j == 0 ? mul(d, d) : add(acc2, mul(d, d));
      }
      guiding = -sqrtf(acc2);
    }
    // everything about the torques is final before the physics: store it now and drop the registers
    st_stream4(io.wheel_torqs + 4 * i, make_float4(wt[0], wt[1], wt[2], wt[3]));
    if (io.torques) st_stream4(io.torques + 4 * i, make_float4(tq[0], tq[1], tq[2], tq[3]));
    if (io.guiding) __stcs(io.guiding + i, guiding);
    if (io.comp0 && i == 0) {   // env-0 scalars behind observed_components :840-867 (pre-physics part)
      float* c = io.comp0;
      c[0] = desired_v; c[1] = current_v; c[2] = desired_w; c[3] = current_w; c[4] = v_delta; c[5] = omega_delta;
      c[9] = tq[0]; c[10] = tq[1]; c[11] = tq[2]; c[12] = tq[3]; c[13] = ac_left; c[14] = ac_right;
      c[15] = a4[0]; c[16] = a4[1]; c[17] = a4[2]; c[18] = a4[3];
      c[19] = g2; c[26] = guiding; c[28] = g1;
    }
    const float r3 = add(add(add(mul(wt[0], wt[0]), mul(wt[1], wt[1])), mul(wt[2], wt[2])), mul(wt[3], wt[3]));
    // calculate_metrics :770 — linear_acc lags a step: it uses the pre-physics projection
    const float lin_acc = cdiv(sub(current_v, tp.z), dv.dt);
    float smooth = 0.f;
    if (kEval) {                                // KAMMA_eval.py:677-685
      float4 ot = ld_stream4(st.old_torques + 4 * i);
      if (zero_old) ot = make_float4(0.f, 0.f, 0.f, 0.f);
      const float d0 = sub(ot.x, wt[0]), d1 = sub(ot.y, wt[1]), d2 = sub(ot.z, wt[2]), d3 = sub(ot.w, wt[3]);
      smooth = dvd(add(add(add(mul(d0, d0), mul(d1, d1)), mul(d2, d2)), mul(d3, d3)), 10.0f);
      st_stream4(st.old_torques + 4 * i, make_float4(wt[0], wt[1], wt[2], wt[3]));
    }
    float4 replay_act = make_float4(0.f, 0.f, 0.f, 0.f);
    if (replay) {   // what record_transition stores as the action (base_KA_DDPG.py:212 relabel, or the policy's)
      if (io.relabel == OIS_RELABEL_EXECUTED)
        replay_act = make_float4(cdiv(wt[0], dv.action_scale), cdiv(wt[1], dv.action_scale), cdiv(wt[2], dv.action_scale),
                                 cdiv(wt[3], dv.action_scale));
      else replay_act = make_float4(a4[0], a4[1], a4[2], a4[3]);
      const int64_t ri = replay_row_base(io, n) + i;
      if (io.relabel == OIS_RELABEL_EXECUTED || p.act_dim == 4) st_stream4(io.replay_actions + 4 * ri, replay_act);
      else if (p.act_dim == 2) reinterpret_cast<float2*>(io.replay_actions)[ri] = make_float2(a4[0], a4[2]);
      else io.replay_actions[ri] = a4[0];
    }

    // ------------------------------------------------------------------ physics :646-653 (replaces PhysX)
    Skid s;
    s.px = pv.x; s.py = pv.y; s.yaw = pv.z; s.vx = pv.w; s.vy = tp.x; s.r = tp.y;
    s.w[0] = wv.x; s.w[1] = wv.y; s.w[2] = wv.z; s.w[3] = wv.w;
    if (p.physics) skid_steps(s, wt, p);

    // ------------------------------------------------------------------ group B: issued one iteration ago
    cp_async_wait<1>();               // pending at most: A(i + stride)
    float4 sa = s_b.sa[tid];
    float4 sb = s_b.sb[tid];
    int64_t eb = s_b.eb[tid];
    float4 se = make_float4(0.f, 0.f, 0.f, 0.f);
    if (kEval) se = s_b.se[tid];
    float prev_obs[6];
    if (replay) {
      const float2 o0 = s_b.obs[0][tid], o1 = s_b.obs[1][tid], o2 = s_b.obs[2][tid];
      prev_obs[0] = o0.x; prev_obs[1] = o0.y; prev_obs[2] = o1.x; prev_obs[3] = o1.y; prev_obs[4] = o2.x; prev_obs[5] = o2.y;
    }
    issue_b(i + stride);              // lands during the rest of this iteration and the next env's physics
    if (kEval) {
      se.x = add(se.x, smooth);
      se.z = add(se.z, dvd(desired_v, 10.0f));
    }

    // ------------------------------------------------------------------ post_physics_step :675-676
    eb += 1;
    sim += 1;
    // is_done :736-760 (has_fallen is identically false for the planar model; env origins are zero :224)
    const bool timeout = eb >= (int64_t)(p.max_episode_length - 1);
    const bool oob = (s.px < -p.bound) | (s.px > p.bound) | (s.py < -p.bound) | (s.py > p.bound);
    const bool reset = timeout | oob;

    // calculate_metrics :767-810
    const float ang_acc = cdiv(sub(s.r, tp.w), dv.dt);
    const float prev_lin = current_v, prev_ang = s.r;
    const float r1 = add(mul(v_delta, v_delta), mul(5.0f, mul(omega_delta, omega_delta)));
    const float r2 = add(mul(lin_acc, lin_acc), mul(ang_acc, ang_acc));
    const float c1 = mul(-25.0f, r1), c2 = mul(-0.02f, r2), c3 = mul(-0.09f, r3);
    const float rdense = add(add(c1, c2), c3);
    const float sparse = (fabsf(v_delta) < 0.01f && fabsf(omega_delta) < 0.01f) ? 0.2f : 0.0f;
    const float observed = add(rdense, sparse);
    const float rew = add(mul(sub(1.0f, g2), observed), mul(g2, guiding));
    sa.x = add(sa.x, c1); sa.y = add(sa.y, c2); sa.z = add(sa.z, c3); sa.w = add(sa.w, rdense);
    sb.x = add(sb.x, sparse); sb.y = add(sb.y, guiding); sb.z = add(sb.z, observed); sb.w = add(sb.w, rew);
    if (kEval) se.y = add(se.y, dvd(fabsf(v_delta), 10.0f));   // KAMMA_eval.py:840

    if (kStats) { acc[19] += g1; acc[20] += g2; }
    if (io.comp0 && i == 0) {   // env-0 scalars, post-physics part
      float* c = io.comp0;
      c[6] = lin_acc; c[7] = ang_acc; c[20] = c1; c[21] = c2; c[22] = c3; c[23] = rdense; c[24] = sparse; c[25] = observed;
      c[27] = rew;
    }

    // ------------------------------------------------------------------ reset_idx :417-498
    if (reset) {
      any_reset_local = 1;
      if (kStats) {   // episode statistics (logging) :470-488
        const float T = p.max_episode_length_s, fin = sb.w, om = sub(1.0f, g2);
        acc[0] += sa.x; acc[1] += sa.y; acc[2] += sa.z; acc[3] += sa.w;
        acc[4] += sb.x; acc[5] += sb.y; acc[6] += sb.z; acc[7] += sb.w;
        acc[8] += se.x; acc[9] += se.y; acc[10] += se.z;
        acc[11] += mul(T, dvd(mul(mul(100.0f, om), sa.x), fin));
        acc[12] += mul(T, dvd(mul(mul(100.0f, om), sa.y), fin));
        acc[13] += mul(T, dvd(mul(mul(100.0f, om), sa.z), fin));
        acc[14] += mul(T, dvd(mul(mul(100.0f, om), sa.w), fin));
        acc[15] += mul(T, dvd(mul(mul(100.0f, om), sb.x), fin));
        acc[16] += mul(T, dvd(mul(mul(100.0f, g2), sb.y), fin));
        acc[17] += mul(T, dvd(mul(mul(100.0f, om), sb.z), fin));
        acc[18] += 1.0f;
      }
      ResetDraws d;
      if (io.reset_uniforms) {
        const float4 ua = __ldcs(reinterpret_cast<const float4*>(io.reset_uniforms + 8 * i));
        const float4 ub = __ldcs(reinterpret_cast<const float4*>(io.reset_uniforms + 8 * i + 4));
        d.u[0] = ua.x; d.u[1] = ua.y; d.u[2] = ua.z; d.u[3] = ua.w; d.u[4] = ub.x; d.u[5] = ub.y; d.u[6] = ub.z; d.u[7] = ub.w;
      } else {   // resets are rare: the first Philox block is recomputed here rather than held across the physics
        d = philox_reset_draws(io.philox_seed, env_philox_offset(io), i, Philox(io.philox_seed)((uint64_t)i, env_philox_offset(io) * 4));
      }
      apply_reset(p, d, sim, s, cmd);
      eb = 0;                                                   // :462
      st.episode_count[i] += 1;                                 // :463
      sa = make_float4(0.f, 0.f, 0.f, 0.f);                     // :485
      sb = sa;
      se = sa;
    }
    if (io.comp0 && i == 0) io.comp0[8] = (float)st.episode_count[0];

    // ------------------------------------------------------------------ get_observations :833 + stores
    const float ob[6] = {cmd.x, cmd.z, v_delta, omega_delta, lin_acc, ang_acc};
    float2* o = reinterpret_cast<float2*>(io.obs + 6 * i);
    o[0] = make_float2(ob[0], ob[1]);
    o[1] = make_float2(ob[2], ob[3]);
    o[2] = make_float2(ob[4], ob[5]);
    float oc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) oc[k] = io.clip_obs > 0.f ? clipf(ob[k], -io.clip_obs, io.clip_obs) : ob[k];
    if (io.obs_clipped) {   // what the env wrapper returns: clamp(obs_buf, +-clipObservations)  (rl_task.py:99)
      float2* q = reinterpret_cast<float2*>(io.obs_clipped + 6 * i);
      q[0] = make_float2(oc[0], oc[1]);
      q[1] = make_float2(oc[2], oc[3]);
      q[2] = make_float2(oc[4], oc[5]);
    }
    __stcs(io.rew + i, rew);
    io.reset[i] = reset ? 1 : 0;
    if (io.timeout) __stcs(io.timeout + i, (int64_t)(timeout ? 1 : 0));

    st_stream4(st.pose_vx + 4 * i, make_float4(s.px, s.py, s.yaw, s.vx));
    st_stream4(st.twist_prev + 4 * i, make_float4(s.vy, s.r, prev_lin, prev_ang));
    st_stream4(st.wheel_speed + 4 * i, make_float4(s.w[0], s.w[1], s.w[2], s.w[3]));
    if (reset) st_stream4(st.commands + 4 * i, cmd);   // only reset_idx changes the commands
    st_stream4(st.sums_a + 4 * i, sa);
    st_stream4(st.sums_b + 4 * i, sb);
    __stcs(st.sim_steps + i, sim);
    __stcs(st.episode_buf + i, eb);
    if (kEval) st_stream4(st.sums_eval + 4 * i, se);

    if (replay) {   // DDPG.record_transition -> memory.add_samples (ddpg.py:313), row `memory_index`
      if (io.clip_obs > 0.f) {   // the trainer's `states` are the wrapper's clamped observations of the previous step
#pragma unroll
        for (int k = 0; k < 6; ++k) prev_obs[k] = clipf(prev_obs[k], -io.clip_obs, io.clip_obs);
      }
      const int64_t ri = replay_row_base(io, n) + i;
      float2* rs = reinterpret_cast<float2*>(io.replay_states + 6 * ri);
      rs[0] = make_float2(prev_obs[0], prev_obs[1]);
      rs[1] = make_float2(prev_obs[2], prev_obs[3]);
      rs[2] = make_float2(prev_obs[4], prev_obs[5]);
      float2* rn = reinterpret_cast<float2*>(io.replay_next_states + 6 * ri);
      rn[0] = make_float2(oc[0], oc[1]);
      rn[1] = make_float2(oc[2], oc[3]);
      rn[2] = make_float2(oc[4], oc[5]);
      io.replay_rewards[ri] = rew;
      io.replay_terminated[ri] = reset ? 1 : 0;
      io.replay_truncated[ri] = 0;
    }
  }

  if (kStats) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int k = 0; k < OIS_EP_STATS_LEN; ++k) {
      const float v = warp_sum(acc[k]);
      if (lane == 0 && v != 0.f) atomicAdd(&s_stats[k], v);
    }
    __syncthreads();
    if (threadIdx.x < OIS_EP_STATS_LEN && s_stats[threadIdx.x] != 0.f) atomicAdd(io.ep_stats + threadIdx.x, s_stats[threadIdx.x]);
  }
  if (kEval && io.any_reset_flags != nullptr) {
    if (__syncthreads_or(any_reset_local) && threadIdx.x == 0) atomicOr(io.any_reset_flags + (env_flag_slot(io) + 1) % 3, 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) io.any_reset_flags[(env_flag_slot(io) + 2) % 3] = 0;
  }
  ENV_STAMP_MAX(3);
  // Graph replay: the last CTA to finish moves the step-control block to the next step.  Every reader of the block — the
  // actor kernel, which completed before this grid's pdl_wait returned, and all CTAs of this grid, which take their
  // ticket only after their last use of it — is done by then.
  if (io.ctl != nullptr && io.ctl->self_advance) {
    __syncthreads();
    if (threadIdx.x == 0) {
      OisStepCtl* c = const_cast<OisStepCtl*>(io.ctl);
      // (the barrier above orders this CTA's stores before the fence; with a host waiting on done_flag they may be stores
      //  to pinned host memory, so the fence is a system-scope one)
      if (io.done_flag) __threadfence_system();
      else __threadfence();
      if (atomicAdd(&c->ticket, 1u) == gridDim.x - 1) {
        if (io.done_flag) {     // every CTA's results are out (each fenced system-wide before its ticket): tell the host which
          __threadfence();      // step is complete — before the block's own update, which the host does not wait for
          *reinterpret_cast<volatile uint32_t*>(io.done_flag) = (uint32_t)(c->step + 1);
        }
        c->ticket = 0;
        step_ctl_advance(c);
      }
    }
  }
}

__global__ void __launch_bounds__(kEnvThreads)
dynamics_kernel(const __grid_constant__ OisEnvParams p, const __grid_constant__ OisEnvState st,
                const float* __restrict__ torques, const int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 pv = ld_stream4(st.pose_vx + 4 * i);
    float4 tp = ld_stream4(st.twist_prev + 4 * i);
    const float4 wv = ld_stream4(st.wheel_speed + 4 * i);
    const float4 t = ld_stream4(torques + 4 * i);
    Skid s;
    s.px = pv.x; s.py = pv.y; s.yaw = pv.z; s.vx = pv.w; s.vy = tp.x; s.r = tp.y;
    s.w[0] = wv.x; s.w[1] = wv.y; s.w[2] = wv.z; s.w[3] = wv.w;
    const float tau[4] = {t.x, t.y, t.z, t.w};
    skid_steps(s, tau, p);
    tp.x = s.vy; tp.y = s.r;
    st_stream4(st.pose_vx + 4 * i, make_float4(s.px, s.py, s.yaw, s.vx));
    st_stream4(st.twist_prev + 4 * i, tp);
    st_stream4(st.wheel_speed + 4 * i, make_float4(s.w[0], s.w[1], s.w[2], s.w[3]));
  }
}

__global__ void __launch_bounds__(kEnvThreads)
reset_idx_kernel(const __grid_constant__ OisEnvParams p, const __grid_constant__ OisEnvState st,
                 const int64_t* __restrict__ ids, const int64_t n_ids, const float* __restrict__ uniforms,
                 const uint64_t seed, const uint64_t offset) {
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n_ids; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = ids[t];
    ResetDraws d;
    if (uniforms) {
      for (int k = 0; k < 8; ++k) d.u[k] = uniforms[8 * i + k];
    } else {
      d = philox_reset_draws(seed, offset, i, Philox(seed)((uint64_t)i, offset * 4));
    }
    Skid s;
    float4 cmd = ld_stream4(st.commands + 4 * i);
    float4 tp = ld_stream4(st.twist_prev + 4 * i);
    apply_reset(p, d, st.sim_steps[i], s, cmd);
    st_stream4(st.pose_vx + 4 * i, make_float4(s.px, s.py, s.yaw, 0.f));
    tp.x = 0.f; tp.y = 0.f;     // previous_* velocities are NOT reset (KAMMA_medium.py:435-446)
    st_stream4(st.twist_prev + 4 * i, tp);
    st_stream4(st.wheel_speed + 4 * i, make_float4(0.f, 0.f, 0.f, 0.f));
    st_stream4(st.commands + 4 * i, cmd);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    st_stream4(st.sums_a + 4 * i, z);
    st_stream4(st.sums_b + 4 * i, z);
    if (p.eval_sums) st_stream4(st.sums_eval + 4 * i, z);
    st.episode_buf[i] = 0;
    st.episode_count[i] += 1;
  }
}

static int grid_for(int64_t n, int threads) {
  int64_t blocks = (n + threads - 1) / threads;
  // waves of resident CTAs (tail balance against per-thread prefetch depth), measured on B200 (profiles/env_waves_sweep.py):
  // 1M envs 3 / 4 / 6 / 8 / 12 / 16 waves = 64.9 / 66.8 / 67.9 / 67.0 / 65.4 / 65.4 % of measured HBM; 4M envs 4 / 8 / 16 =
  // 68.3 / 71.5 / 73.0 %.  Block sizes 64 / 96 / 256 at equal warps per SM change nothing.
  static const int waves_env = getenv("OIS_ENV_WAVES") ? atoi(getenv("OIS_ENV_WAVES")) : 0;
  const int waves = waves_env > 0 ? waves_env : (n >= ((int64_t)1 << 21) ? 16 : 8);
  const int64_t cap = (int64_t)num_sms() * kEnvMinBlocks * waves;   // `waves` waves of resident CTAs; grid-stride beyond
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// Is q = RN(q0 + (x - d q0) rc), q0 = RN(x rc), rc = RN(1/d), the correctly rounded x / d for EVERY x?  The answer
// depends only on the significands, so all 2^23 of x in [1, 2) are tried once per divisor (about 20 ms on one host core;
// cached).  Mirrors cdiv() with the host's correctly rounded fmaf.
static bool const_div_exact(float d) {
  static std::mutex mu;
  static std::map<uint32_t, bool> cache;
  uint32_t key;
  memcpy(&key, &d, 4);
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  bool ok = d > 1e-12f && d < 1e12f;       // comfortably inside the normal range: no under- / overflow in q0, r
  if (ok) {
    const volatile float rcv = 1.0f / d;   // IEEE division: RN(1/d)
    const float rc = rcv;
    for (uint32_t m = 0; m < (1u << 23) && ok; ++m) {
      const uint32_t bits = 0x3F800000u | m;
      float x;
      memcpy(&x, &bits, 4);
      const volatile float q0v = x * rc;
      const float q0 = q0v;
      const float r = fmaf(-d, q0, x);
      const float q = fmaf(r, rc, q0);
      const volatile float refv = x / d;
      if (q != refv) ok = false;
    }
  }
  cache[key] = ok;
  return ok;
}

static ConstDiv const_div(float d) {
  ConstDiv c;
  c.d = d;
  c.exact = const_div_exact(d) ? 1 : 0;
  const volatile float rc = 1.0f / d;
  c.rc = rc;
  return c;
}

EnvDerived make_env_derived(const OisEnvParams* params) {
  EnvDerived dv;
  dv.dt = const_div(params->dt);
  dv.action_scale = const_div(params->action_scale);
  dv.max_sim_steps = const_div(params->max_sim_steps);
  return dv;
}

static int check_state(const OisEnvParams* p, const OisEnvState* s) {
  OIS_CHECK_ARG(p && s, "null params/state");
  OIS_CHECK_ARG(p->act_dim == 1 || p->act_dim == 2 || p->act_dim == 4, "act_dim must be 1, 2 or 4");
  OIS_CHECK_ARG(s->pose_vx && s->twist_prev && s->wheel_speed && s->commands && s->sums_a && s->sums_b,
                "null state array");
  OIS_CHECK_ARG(s->sim_steps && s->episode_buf && s->episode_count, "null counter array");
  OIS_CHECK_ARG(!p->eval_sums || (s->sums_eval && s->old_torques), "eval_sums needs sums_eval and old_torques");
  OIS_CHECK_ARG(p->decimation >= 0 && p->decimation <= 1000, "bad decimation");
  return OIS_OK;
}

}  // namespace ois

using namespace ois;

extern "C" int ois_env_step(const OisEnvParams* params, const OisEnvState* state, const OisEnvIO* io, int64_t n_envs,
                            void* stream) {
  if (int e = check_state(params, state)) return e;
  OIS_CHECK_ARG(io, "null io");
  OIS_CHECK_ARG(n_envs >= 0, "negative n_envs");
  OIS_CHECK_ARG(io->actions && io->obs && io->rew && io->reset && io->wheel_torqs, "null required io array");
  const bool any_replay = io->replay_states || io->replay_actions || io->replay_rewards || io->replay_next_states ||
                          io->replay_terminated || io->replay_truncated;
  const bool all_replay = io->replay_states && io->replay_actions && io->replay_rewards && io->replay_next_states &&
                          io->replay_terminated && io->replay_truncated;
  OIS_CHECK_ARG(!any_replay || all_replay, "replay pointers must be all set or all null");
  OIS_CHECK_ARG(io->flag_slot >= 0 && io->flag_slot < 3, "flag_slot must be 0..2");
  OIS_CHECK_ARG(!io->done_flag || io->ctl, "done_flag needs a step-control block (ctl) with self_advance");
  if (n_envs == 0) return OIS_OK;
  const int grid = grid_for(n_envs, kEnvThreads);
  cudaStream_t s = (cudaStream_t)stream;
  {   // an actor launch deferred by ois_agent_act on this thread: fuse with it, or let it go first (agent.cu)
    const int f = rollout_try_fuse(params, state, io, n_envs, s);
    if (f < 0) return f;
    if (f == 1) return OIS_OK;
  }
  static bool carveout_set = false;   // 32 KB of staging per CTA x kEnvMinBlocks CTAs: ask for the large shared-memory split
  if (!carveout_set) {
    cudaFuncSetAttribute(env_step_kernel<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(env_step_kernel<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(env_step_kernel<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    cudaFuncSetAttribute(env_step_kernel<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    carveout_set = true;
  }
  const bool stats = io->ep_stats != nullptr, ev = params->eval_sums != 0;
  const EnvDerived dv = make_env_derived(params);
  cudaError_t le;
  if (stats && ev) le = launch_pdl(env_step_kernel<true, true>, grid, kEnvThreads, 0, s, *params, *state, *io, n_envs, dv);
  else if (stats) le = launch_pdl(env_step_kernel<true, false>, grid, kEnvThreads, 0, s, *params, *state, *io, n_envs, dv);
  else if (ev) le = launch_pdl(env_step_kernel<false, true>, grid, kEnvThreads, 0, s, *params, *state, *io, n_envs, dv);
  else le = launch_pdl(env_step_kernel<false, false>, grid, kEnvThreads, 0, s, *params, *state, *io, n_envs, dv);
  OIS_CHECK_CUDA(le);
  return OIS_OK;
}

extern "C" int ois_debug_env_trace(unsigned long long* out_host, int reset) {
  if (reset) {
    const unsigned long long z[4] = {~0ull, ~0ull, 0ull, 0ull};
    return cudaMemcpyToSymbol(ois::g_env_trace, z, sizeof(z)) == cudaSuccess ? 0 : -2;
  }
  return cudaMemcpyFromSymbol(out_host, ois::g_env_trace, sizeof(unsigned long long) * 4) == cudaSuccess ? 0 : -2;
}

extern "C" int ois_dynamics_step(const OisEnvParams* params, const OisEnvState* state, const float* wheel_torqs,
                                 int64_t n_envs, void* stream) {
  OIS_CHECK_ARG(params && state && wheel_torqs, "null argument");
  OIS_CHECK_ARG(state->pose_vx && state->twist_prev && state->wheel_speed, "null state array");
  OIS_CHECK_ARG(n_envs >= 0, "negative n_envs");
  if (n_envs == 0) return OIS_OK;
  dynamics_kernel<<<grid_for(n_envs, kEnvThreads), kEnvThreads, 0, (cudaStream_t)stream>>>(*params, *state, wheel_torqs, n_envs);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

extern "C" int ois_env_reset_idx(const OisEnvParams* params, const OisEnvState* state, const int64_t* env_ids,
                                 int64_t n_ids, const float* reset_uniforms, uint64_t philox_seed,
                                 uint64_t philox_offset, int64_t n_envs, void* stream) {
  if (int e = check_state(params, state)) return e;
  OIS_CHECK_ARG(n_ids >= 0 && n_ids <= n_envs, "bad n_ids");
  OIS_CHECK_ARG(env_ids || n_ids == 0, "null env_ids");
  if (n_ids == 0) return OIS_OK;
  reset_idx_kernel<<<grid_for(n_ids, kEnvThreads), kEnvThreads, 0, (cudaStream_t)stream>>>(
      *params, *state, env_ids, n_ids, reset_uniforms, philox_seed, philox_offset);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}
