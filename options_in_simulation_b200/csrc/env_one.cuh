// ONE env's policy step as a straight-line device function: the loop body of env_step_kernel<false, false> (env_step.cu)
// with its cp.async staging replaced by values the caller loaded beforehand (EnvLoaded).  It exists for the fused policy
// step of mlp_tct.cu, where the warp that has a tile's actions steps those envs itself; the stand-alone kernel keeps its
// own copy of this text because its staging is interleaved with the arithmetic.  The two must stay the same arithmetic,
// statement for statement: tests/test_rollout_graph_gpu.py holds the fused step bit-identical to act + env step.
#pragma once
#ifndef ENV_ONE_STAMP
#define ENV_ONE_STAMP(slot) do { } while (0)   // timeline diagnostics of the including kernel (mlp_tct.cu)
#endif
#include "env_core.cuh"

namespace ois {

struct EnvLoaded {
  float4 pv, tp, wv, cmd, act, sa, sb;
  int64_t sim, eb;
  float2 obs0, obs1, obs2;     // the previous raw observation (the replay row's `states`), read only when replay is fused
};

// everything of env i but its action
__device__ __forceinline__ void env_load(const OisEnvState& st, const OisEnvIO& io, int64_t i, EnvLoaded& L) {
  L.pv = ld_stream4(st.pose_vx + 4 * i);
  L.tp = ld_stream4(st.twist_prev + 4 * i);
  L.wv = ld_stream4(st.wheel_speed + 4 * i);
  L.cmd = ld_stream4(st.commands + 4 * i);
  L.sa = ld_stream4(st.sums_a + 4 * i);
  L.sb = ld_stream4(st.sums_b + 4 * i);
  L.sim = st.sim_steps[i];
  L.eb = st.episode_buf[i];
  L.obs0 = L.obs1 = L.obs2 = make_float2(0.f, 0.f);
  if (io.replay_states != nullptr) {
    const float2* o = reinterpret_cast<const float2*>(io.obs + 6 * i);
    L.obs0 = o[0]; L.obs1 = o[1]; L.obs2 = o[2];
  }
}

// returns 1 when the env was reset this step
// `act_fn()` yields the env's action row; it is called as late as the arithmetic allows, so that a caller whose actions
// arrive late (it may block in there) has everything that depends on the state alone behind it by then.
template <class ActFn>
__device__ __forceinline__ int env_one_direct(const OisEnvParams& p, const OisEnvState& st, const OisEnvIO& io, const EnvDerived& dv,
                                              const int64_t n, const int64_t i, const EnvLoaded& L, ActFn&& act_fn,
                                              const bool defer_obs_clipped, float (&obs_clipped_out)[6]) {
  constexpr bool kStats = false, kEval = false;
  float acc[OIS_EP_STATS_LEN] = {};     // dead (kStats is false): keeps the shared text compiling
  int any_reset_local = 0;
  const bool zero_old = false;
  const bool replay = io.replay_states != nullptr;
  // ------------------------------------------------------------------ group A: this env's now, the next one's in flight
  const float4 pv = L.pv;
  const float4 tp = L.tp;
  const float4 wv = L.wv;
  float4 cmd = L.cmd;
  int64_t sim = L.sim;
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
  float mask = 0.f, nmask = 0.f, ng = 0.f;
  if (p.blend == OIS_BLEND_SWITCH) {          // :632-636
    const float rm = need_mask_draw ? u01(philox2x32((uint32_t)i, (uint32_t)env_philox_offset(io),
                                                     (uint32_t)io.philox_seed ^ (uint32_t)(io.philox_seed >> 32) ^ (uint32_t)(i >> 32)))
                                    : __ldcs(io.rand_mask + i);
    mask = rm < g1 ? 1.0f : 0.0f;
    nmask = sub(1.0f, mask);
  } else {                                    // KA_DDPG_1D.py:632-633
    ng = sub(1.0f, g1);
  }
  // calculate_metrics :770 — linear_acc lags a step: it uses the pre-physics projection
  const float lin_acc = cdiv(sub(current_v, tp.z), dv.dt);
  SkidPre skid_pre = {};
  if (p.physics) skid_pre = skid_prepare(pv.z, p);
  // ------------------------------------------------------------------ the actions (everything above needs the state alone)
  float a4[4];
  {
    const float4 a = act_fn();
    if (p.act_dim == 4) { a4[0] = a.x; a4[1] = a.y; a4[2] = a.z; a4[3] = a.w; }            // KAMMA_medium.py:600
    else if (p.act_dim == 2) { a4[0] = a.x; a4[1] = a.x; a4[2] = a.y; a4[3] = a.y; }       // KA_DDPG_2D.py:624
    else { a4[0] = a4[1] = a4[2] = a4[3] = a.x; }                                          // KA_DDPG_1D.py:600
  }
  if (io.clip_actions > 0.f) {      // VecEnvRLGames.step clamps the actions before pre_physics_step (clipActions)
#pragma unroll
    for (int j = 0; j < 4; ++j) a4[j] = clipf(a4[j], -io.clip_actions, io.clip_actions);
  }
  if (p.blend == OIS_BLEND_SWITCH) {
#pragma unroll
    for (int j = 0; j < 4; ++j) tq[j] = add(mul(mask, crit[j]), mul(nmask, mul(a4[j], p.action_scale)));
  } else {
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
  ENV_ONE_STAMP(31);
  if (p.physics) skid_steps(s, wt, p, skid_pre);
  ENV_ONE_STAMP(32);

  // ------------------------------------------------------------------ group B: issued one iteration ago
  float4 sa = L.sa;
  float4 sb = L.sb;
  int64_t eb = L.eb;
  float4 se = make_float4(0.f, 0.f, 0.f, 0.f);
  if (kEval) se = make_float4(0.f, 0.f, 0.f, 0.f);
  float prev_obs[6];
  if (replay) {
    const float2 o0 = L.obs0, o1 = L.obs1, o2 = L.obs2;
    prev_obs[0] = o0.x; prev_obs[1] = o0.y; prev_obs[2] = o1.x; prev_obs[3] = o1.y; prev_obs[4] = o2.x; prev_obs[5] = o2.y;
  }
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
#pragma unroll
  for (int k = 0; k < 6; ++k) obs_clipped_out[k] = oc[k];
  if (defer_obs_clipped) {   // the caller stores the tile's rows itself (coalesced: mlp_tct.cu)
  } else if (io.obs_clipped) {   // what the env wrapper returns: clamp(obs_buf, +-clipObservations)  (rl_task.py:99)
    float2* q = reinterpret_cast<float2*>(io.obs_clipped + 6 * i);
    q[0] = make_float2(oc[0], oc[1]);
    q[1] = make_float2(oc[2], oc[3]);
    q[2] = make_float2(oc[4], oc[5]);
  }
  __stcs(io.rew + i, rew);
  ENV_ONE_STAMP(33);
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
  (void)acc;
  return any_reset_local;
}

}  // namespace ois
