/*
 * ois_b200.h — C ABI of libois_b200.so: the B200-native (sm_100a) rollout-and-update hot path of
 * meesjansen/Options_In_Simulation.
 *
 * The reference has no FFI/plugin interface (it is 100 % Python on top of Isaac Sim + skrl); the
 * drop-in boundary is its two duck-typed Python APIs (RLTask and skrl Agent).  The Python mirrors
 * in options_in_simulation_b200/{envs,agent,memory}.py keep those names and call the entry points
 * below through ctypes.  Each entry point cites the reference code it replaces
 * (paths relative to the reference repository root).
 *
 * Conventions
 *   - every function returns 0 on success or a negative OIS_ERR_* code; nothing throws across the ABI;
 *     ois_last_error_string() describes the last failure on the calling thread
 *   - all data pointers are DEVICE pointers unless the name ends in _host; buffers are caller-allocated,
 *     the callee never frees or retains them beyond the call (handles excepted)
 *   - launches are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy stream);
 *     no host synchronisation inside any call
 *   - [N,4] float arrays must be 16-byte aligned, obs [N,6] 8-byte aligned
 *   - integers that are torch.long at the reference's Python surface are int64_t here
 */
#ifndef OIS_B200_H
#define OIS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OIS_OK 0
#define OIS_ERR_INVALID_ARG (-1)
#define OIS_ERR_CUDA (-2)
#define OIS_ERR_UNSUPPORTED (-3)

#define OIS_ABI_VERSION 3

/* ------------------------------------------------------------------------------------------------
 * Environment: TorqueDistributionTask (src/my_envs/KAMMA_medium.py and its variants)
 * ---------------------------------------------------------------------------------------------- */

enum { OIS_BLEND_LINEAR = 0, OIS_BLEND_SWITCH = 1 };          /* KA_DDPG_1D.py:633 | KAMMA_medium.py:632-636 */
enum { OIS_GAMMA_CONST = 0, OIS_GAMMA_DECAY = 1 };            /* KAMMA_medium.py:627-628 */
enum { OIS_CMD_RAND = 0, OIS_CMD_GV = 1, OIS_CMD_EVAL = 2, OIS_CMD_BD = 3 };
/* KAMMA_medium.py:569-575 | KAMMA_curr1.py:577-599 (box, "gv") | KAMMA_eval.py:545-550 | KAMMA_curr2.py:601-641 (bimodal, "bd") */
enum { OIS_RELABEL_NONE = 0, OIS_RELABEL_EXECUTED = 1 };      /* base_KA.py | base_KA_DDPG.py:212 */

typedef struct OisEnvParams {
  int32_t act_dim;            /* 1 | 2 | 4   (KA_DDPG_1D.py:600, KA_DDPG_2D.py:624, KAMMA_medium.py:600) */
  int32_t blend;              /* OIS_BLEND_* */
  int32_t gamma1_mode;        /* action assist */
  int32_t gamma2_mode;        /* reward assist */
  int32_t guiding_post_clip;  /* 0: -||a*scale - criteria|| (KA_DDPG_1D.py:642); 1: -||clip(torques) - criteria|| (KAMMA_medium.py:656) */
  int32_t command_mode;       /* OIS_CMD_* */
  int32_t physics;            /* 0: state untouched (parity tests), 1: skid-steer model (replaces PhysX, KAMMA_medium.py:646-653) */
  int32_t eval_sums;          /* 1: Smoothness / Tracking error / Desired velocity sums (KAMMA_eval.py:677-685,840) */
  int32_t decimation;         /* KAMMA_medium.py:75 */
  int32_t max_episode_length; /* KAMMA_medium.py:262 */
  float gamma1_const, gamma2_const;
  float max_sim_steps;        /* KAMMA_medium.py:179 */
  float kp, kp_omega, vehicle_mass, vehicle_inertia, dt; /* KAMMA_medium.py:263,613,174,175,260 */
  float action_scale, torque_clip, bound;                 /* KAMMA_medium.py:238,649,187 */
  float x_lo, x_range, y_lo, y_range, h_lo, h_range;      /* command ranges: lower and (upper-lower), KAMMA_medium.py:245-247 */
  float x_hi;
  float max_episode_length_s;                             /* KAMMA_medium.py:261 */
  /* skid-steer model (oracle/dynamics_twin.py states the equations) */
  float mass, yaw_inertia, wheel_inertia, wheel_radius, half_wheelbase, half_track;
  float mu_n;                 /* mu * m * g / 4 */
  float slip_eps, wheel_damping, wheel_speed_max, sim_dt;
} OisEnvParams;

/* Per-env state, struct of [N,4] float arrays + int64 counters. */
typedef struct OisEnvState {
  float* pose_vx;        /* [N,4] px, py, yaw, vx      (vx == v_forward_projected, KAMMA_medium.py:710) */
  float* twist_prev;     /* [N,4] vy, wz, previous_linear_velocity[:,0], previous_angular_velocity[:,2] */
  float* wheel_speed;    /* [N,4] dof_vel, order fl, rl, fr, rr (KAMMA_medium.py:623) */
  float* commands;       /* [N,4] x, y, yaw-rate, heading (KAMMA_medium.py:378-380) */
  float* sums_a;         /* [N,4] episode_sums r1, r2, r3, Dense          (KAMMA_medium.py:803-806) */
  float* sums_b;         /* [N,4] episode_sums Sparse, Guiding, Observed, Final (KAMMA_medium.py:807-810) */
  float* sums_eval;      /* [N,4] Smoothness, Tracking error, Desired velocity, 0 — eval_sums only, else NULL */
  float* old_torques;    /* [N,4] eval_sums only, else NULL (KAMMA_eval.py:205) */
  int64_t* sim_steps;    /* [N] */
  int64_t* episode_buf;  /* [N] */
  int64_t* episode_count;/* [N] */
} OisEnvState;

/* Device-resident per-step scalars, for CUDA-graph replay of the rollout step (the graph's kernel arguments are frozen
 * at capture, so what changes from one policy step to the next lives here instead).  When OisEnvIO.ctl / OisActIO.ctl is
 * non-NULL the kernels take these values in place of the like-named by-value fields.  The block moves to the next step
 * either inside ois_env_step (self_advance = 1: the last CTA of the env kernel to finish does it, after every reader of
 * the block is done) or with ONE single-thread launch of ois_step_ctl_advance at the end of the captured step.  The host initialises it with
 * a plain host->device copy of this struct. */
typedef struct OisStepCtl {
  int64_t step;                /* steps taken since the host initialised the block */
  uint64_t env_philox_offset;  /* OisEnvIO.philox_offset of this step */
  uint64_t act_philox_offset;  /* OisActIO.philox_offset of this step */
  int64_t replay_row;          /* memory_index: the ring slot the fused append writes (ddpg.py:313 -> Memory.add_samples) */
  int64_t replay_rows;         /* memory_size */
  int32_t flag_slot;           /* OisEnvIO.flag_slot of this step */
  float noise_scale;           /* OisActIO.noise_scale of this step (ddpg.py:247-251) */
  double timestep;             /* trainer timestep of this step */
  double expl_timesteps, expl_initial_scale, expl_final_scale;   /* ddpg.py:105-107 */
  int32_t self_advance;        /* 1: ois_env_step advances this block itself (its last CTA to finish, see `ticket`), so a
                                  captured step needs no ois_step_ctl_advance launch */
  uint32_t ticket;             /* CTAs of the current env step that have finished; 0 between steps */
} OisStepCtl;

/* ctl := the next policy step (offsets + 1, replay_row + 1 mod replay_rows, flag_slot + 1 mod 3, timestep + 1 and the
 * decayed noise scale of ddpg.py:247-251 evaluated in double as the host path does). */
int ois_step_ctl_advance(OisStepCtl* ctl, void* stream);

#define OIS_EP_STATS_LEN 24   /* see ois_env_step */
#define OIS_COMP0_LEN 32

typedef struct OisEnvIO {
  /* inputs */
  const float* actions;         /* [N,act_dim] */
  const float* rand_mask;       /* [N] uniforms of KAMMA_medium.py:633, or NULL -> in-kernel Philox */
  const float* reset_uniforms;  /* [N,8] draws of reset_idx in reference order, or NULL -> Philox.  Column 6 is the uniform of
                                   sample_velocity_command (OIS_CMD_BD: the mode choice `rand < 0.5`), column 7 is unused
                                   except for OIS_CMD_BD, where it holds the STANDARD NORMAL behind torch.normal */
  uint64_t philox_seed, philox_offset;
  /* outputs */
  float* obs;                   /* [N,6] in/out (raw obs_buf): read as the previous observation when replay_states != NULL */
  float* rew;                   /* [N]   rew_buf */
  uint8_t* reset;               /* [N]   reset_buf (bool) */
  float* wheel_torqs;           /* [N,4] clipped torques */
  int64_t* timeout;             /* [N]   timeout_buf, nullable */
  float* torques;               /* [N,4] pre-clip execution action, nullable */
  float* guiding;               /* [N]   guiding_reward, nullable */
  float* ep_stats;              /* [OIS_EP_STATS_LEN] zeroed by the caller, nullable: sums over reset envs of
                                   the 8 (+3) episode sums [0..10], of the 7 ratio keys [11..17], count [18],
                                   sum of gamma1 / gamma2 over ALL envs [19],[20] (KAMMA_medium.py:470-488) */
  float* comp0;                 /* [OIS_COMP0_LEN] env-0 scalars behind observed_components, nullable (KAMMA_medium.py:840-867) */
  int32_t* any_reset_flags;     /* [3] eval_sums only: rotating "some env reset last step" flags (KAMMA_eval.py:470) */
  int32_t flag_slot;            /* step index mod 3 */
  /* fused replay append (row `memory_index` of each (memory_size, N, dim) tensor), all nullable together */
  float* replay_states;         /* [N,6]  <- previous obs */
  float* replay_actions;        /* [N,replay_act_dim] <- wheel_torqs/action_scale (RELABEL_EXECUTED, 4 wide) or actions */
  float* replay_rewards;        /* [N] */
  float* replay_next_states;    /* [N,6] */
  uint8_t* replay_terminated;   /* [N] */
  uint8_t* replay_truncated;    /* [N] (always 0: the wrapper finds no time_outs) */
  int32_t relabel;              /* OIS_RELABEL_* */
  float clip_obs;               /* > 0: +-clamp applied to obs_clipped and to the replay rows, as the env wrapper does
                                   (clipObservations, KAMMA_medium.py:42, rl_task.py:99); <= 0: no clamp */
  float clip_actions;           /* > 0: actions are clamped to +-clip_actions before use, as VecEnvRLGames.step does
                                   (clipActions, rl_task.py:100); <= 0: no clamp */
  float* obs_clipped;           /* [N,6] nullable: clamp(obs, +-clip_obs) — what the env wrapper hands to the agent;
                                   `obs` itself stays the raw obs_buf */
  const OisStepCtl* ctl;        /* nullable; when set: philox_offset and flag_slot come from it, and the replay_* pointers
                                   are the BASE of the (memory_size, N, dim) tensors, row ctl->replay_row is written */
  uint32_t* done_flag;          /* nullable, needs ctl with self_advance: a word of PINNED HOST memory (device-visible
                                   address).  When every store of this launch has been performed at system scope — obs /
                                   rew / reset may themselves be pinned host buffers — the kernel's last CTA stores
                                   (uint32_t)ctl->step (the value AFTER the advance, i.e. steps completed) there, so a host
                                   in the loop can wait for the step with ois_host_wait instead of a device synchronise */
} OisEnvIO;

/* One policy step for all envs: pre_physics_step + decimation physics sub-steps + post_physics_step
 * (is_done, calculate_metrics, reset_idx, get_observations), fused.
 * Replaces src/my_envs/KAMMA_medium.py:596-727 (+ variants) and, when the replay pointers are set,
 * DDPG.record_transition's memory.add_samples (src/my_agents/ddpg.py:313). */
int ois_env_step(const OisEnvParams* params, const OisEnvState* state, const OisEnvIO* io, int64_t n_envs,
                 void* stream);

/* Only the dynamics: `decimation` sub-steps of the skid-steer model under constant wheel torques [N,4].
 * Stands where SimulationContext.step is called (src/my_envs/KAMMA_medium.py:646-653). */
int ois_dynamics_step(const OisEnvParams* params, const OisEnvState* state, const float* wheel_torqs,
                      int64_t n_envs, void* stream);

/* reset_idx for an explicit index list (src/my_envs/KAMMA_medium.py:417-498), used by post_reset. */
int ois_env_reset_idx(const OisEnvParams* params, const OisEnvState* state, const int64_t* env_ids, int64_t n_ids,
                      const float* reset_uniforms /* [N,8] indexed by env id, or NULL */, uint64_t philox_seed,
                      uint64_t philox_offset, int64_t n_envs, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Replay memory: skrl RandomMemory / inline FIFOMemory (train/train_kaddpg_1d_nofifo_random_RLIL.py:105-206)
 * ---------------------------------------------------------------------------------------------- */

/* memory.add_samples for one step: copy N rows into ring slot `memory_index` of each tensor. */
int ois_replay_append(float* mem_states, float* mem_actions, float* mem_rewards, float* mem_next_states,
                      uint8_t* mem_terminated, uint8_t* mem_truncated, int64_t memory_index, int64_t n_envs,
                      int32_t obs_dim, int32_t act_dim, const float* states, const float* actions,
                      const float* rewards, const float* next_states, const uint8_t* terminated,
                      const uint8_t* truncated, void* stream);

/* memory.sample_by_index: gather rows idx[b] of the flattened (memory_size*N, dim) views. */
int ois_replay_gather(const float* mem_states, const float* mem_actions, const float* mem_rewards,
                      const float* mem_next_states, const uint8_t* mem_terminated, const uint8_t* mem_truncated,
                      const int64_t* idx, int64_t batch, int64_t mem_rows, int32_t obs_dim, int32_t act_dim,
                      float* out_states, float* out_actions, float* out_rewards, float* out_next_states,
                      uint8_t* out_terminated, uint8_t* out_truncated, void* stream);

/* torch.randperm(len)[:batch] replacement for throughput mode: `batch` distinct indices in [0, len)
 * drawn on the device (Philox + cycle-walking Feistel permutation), no host round trip. */
int ois_replay_sample_indices(int64_t* idx, int64_t batch, int64_t len, uint64_t philox_seed,
                              uint64_t philox_offset, void* stream);

/* Device-resident scalars of a CUDA-graph-captured DDPG._update: what changes from one update to the next.  The host
 * writes the block (plain host->device copy) when its mirrors move for another reason (the memory grew, an eager
 * sample was drawn, a checkpoint was loaded); ois_replay_sample_gather advances it once per launch. */
typedef struct OisSampleCtl {
  int64_t len;         /* len(memory): transitions to sample from (Memory.__len__) */
  uint64_t draws;      /* philox_offset of the next ois_replay_sample_indices-equivalent draw */
  int64_t adam_step;   /* optimizer steps completed; after the launch: the 1-based step of the update it feeds
                          (point OisUpdateIO.adam_step_dev here) */
  uint32_t ticket;     /* CTAs of the running launch that have finished; 0 between launches */
  uint32_t pad_;
} OisSampleCtl;

/* ctl := (len, draws, adam_step, ticket 0), stream-ordered (one single-thread launch; the values are launch arguments). */
int ois_sample_ctl_set(OisSampleCtl* ctl, int64_t len, uint64_t draws, int64_t adam_step, void* stream);

/* ois_replay_sample_indices(len = ctl->len, offset = ctl->draws) and ois_replay_gather of those indices in ONE launch
 * (Memory.sample, train/train_kaddpg_1d_nofifo_random_RLIL.py:130-160); idx also receives the indices; the last CTA
 * advances ctl (draws + 1, adam_step + 1).  Precondition: batch <= ctl->len. */
int ois_replay_sample_gather(const float* mem_states, const float* mem_actions, const float* mem_rewards,
                             const float* mem_next_states, const uint8_t* mem_terminated, const uint8_t* mem_truncated,
                             int64_t* idx, int64_t batch, int32_t obs_dim, int32_t act_dim, float* out_states,
                             float* out_actions, float* out_rewards, float* out_next_states, uint8_t* out_terminated,
                             uint8_t* out_truncated, uint64_t philox_seed, OisSampleCtl* ctl, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Agent: DDPG (src/my_agents/ddpg.py) with the models of train/train_kamma_4d_nofifo_random_RLIL.py:21-52
 * ---------------------------------------------------------------------------------------------- */

#define OIS_MLP_LAYERS 4
enum { OIS_ACT_NONE = 0, OIS_ACT_SIGMOID = 1, OIS_ACT_TANH = 2 };
enum { OIS_PRECISION_FP32 = 0, OIS_PRECISION_BF16 = 1 };

/* 4-layer MLP in->512->512->128->out with ELU between layers; parameters are ONE flat fp32 buffer in
 * nn.Sequential order: W0[h0,in] b0[h0] W1[h1,h0] b1[h1] W2[h2,h1] b2[h2] W3[out,h2] b3[out]. */
typedef struct OisMlpShape {
  int32_t in_dim, h0, h1, h2, out_dim;
  int32_t out_act;  /* OIS_ACT_* : Sigmoid for the training actor, Tanh in eval scripts, none for the critic */
} OisMlpShape;

typedef struct OisScaler {  /* skrl RunningStandardScaler state (fp64), restated in oracle/ref_harness/skrl_shim.py */
  double* running_mean;     /* [obs_dim] */
  double* running_variance; /* [obs_dim] */
  double* current_count;    /* [1] */
} OisScaler;

/* RunningStandardScaler.forward(x, train) on [n_rows, dim] (dim <= 16): y = clamp((x - mean) / (sqrt(var) + 1e-8), +-5);
 * train != 0 first merges the batch statistics into the fp64 running state (parallel variance), as skrl does
 * inside DDPG._update (src/my_agents/ddpg.py:381-382).  Restated in oracle/ref_harness/skrl_shim.py. */
int ois_scaler_forward(const OisScaler* scaler, const float* x, float* y, int64_t n_rows, int32_t dim, int32_t train,
                       void* stream);

typedef struct OisActIO {
  const float* states;       /* [N,obs_dim] */
  float* actions;            /* [N,act_dim] out */
  float* ou_state;           /* [N,act_dim] in/out OU noise state, nullable (no exploration noise) */
  const float* normals;      /* [N,act_dim] N(0,1) draws, or NULL -> Philox */
  const float* uniforms;     /* [N,act_dim] U(0,1) draws for the random_timesteps branch, or NULL -> Philox */
  uint64_t philox_seed, philox_offset;
  int32_t random_phase;      /* 1: timestep < random_timesteps -> U[low,high] (ddpg.py:229-230) */
  float theta, sigma, base_scale; /* OrnsteinUhlenbeckNoise */
  float noise_scale;         /* decayed scale of ddpg.py:247-251, <0: timestep > exploration timesteps (no noise added) */
  float clip_low, clip_high; /* action space bounds (ddpg.py:255) */
  float* noise_stats;        /* [3] max, min, sum of the applied noises (ddpg.py:258-260), nullable */
  const OisStepCtl* ctl;     /* nullable; when set: philox_offset and noise_scale come from it */
} OisActIO;

typedef struct OisAgent OisAgent;   /* opaque: workspaces, bf16 weight copies, TMA descriptors, graphs */

int ois_agent_create(OisAgent** out, const OisMlpShape* actor, const OisMlpShape* critic, int64_t max_rows,
                     int32_t batch_size, int32_t precision, void* stream);
int ois_agent_destroy(OisAgent* agent);
/* The caller changed parameter buffers behind the handle's back (load_state_dict, checkpoint load):
 * drop cached bf16 weight images.  ois_ddpg_update_phase invalidates by itself. */
int ois_agent_invalidate(OisAgent* agent);

/* Bring the bf16 image of the actor weights that ois_agent_act reads for `n_rows` envs up to date with `actor_params`
 * (a no-op when nothing changed — after ois_ddpg_update_phase the image is already current, the Adam kernels write it
 * element by element — and for fp32 handles).  ois_agent_act does this by itself; a captured CUDA graph of the rollout
 * step cannot, so its owner calls this before each replay (DDPG._update changes the weights between replays). */
int ois_agent_sync_weights(OisAgent* agent, const float* actor_params, int64_t n_rows, void* stream);

/* DDPG.act (ddpg.py:215-268): scaler -> actor MLP -> OU noise -> clamp, N rows. */
int ois_agent_act(OisAgent* agent, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                  int64_t n_rows, void* stream);

/* Rollout fusion.  Between _begin and _end (per host thread) an ois_agent_act that the one-launch policy step can serve
 * (bf16 handle, at most 32 x SMs rows, act_dim <= 4, not the random phase) is recorded instead of launched, and the
 * ois_env_step that follows on the same stream and consumes its `actions` buffer launches actor + env step as ONE kernel
 * (the loop body of src/my_trainers/base_KA_DDPG.py:175-219).  Results are bit-identical to the two launches.  Any other
 * call pattern simply launches the recorded actor first; _end launches it if nothing consumed it. */
int ois_rollout_fuse_begin(void);
int ois_rollout_fuse_end(void);
/* one-launch policy steps issued by this process so far (a caller that counts kernel launches subtracts the difference) */
int64_t ois_rollout_fused_launches(void);

/* Plain MLP forward (Model.act): y[N,out] = mlp(x[N,in]); x already preprocessed. */
int ois_mlp_forward(OisAgent* agent, const OisMlpShape* shape, const float* params, const float* x, float* y,
                    int64_t n_rows, void* stream);

/* ois_mlp_forward without a handle and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same kernels and bit-identical results as ois_mlp_forward (fp32); the hidden activations
 * live in ONE file-scope device array inside the library (csrc/no_alloc.cu: 4096 x (512 + 512 + 128) fp32 = 18 MiB per
 * GPU that loads the library, held until exit) instead of a handle's cudaMalloc'ed workspace.  The call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine).  Limits: n_rows <= OIS_STATIC_MLP_MAX_ROWS and hidden
 * sizes <= 512, 512, 128, else OIS_ERR_INVALID_ARG with ois_last_error_string() naming the limit.  Every call on a device
 * shares the array (it keeps the last call's contents; nothing in it needs to start at zero): calls that overlap on the
 * GPU race, so serialise them on one stream or with events. */
#define OIS_STATIC_MLP_MAX_ROWS 4096
int ois_mlp_forward_static(const OisMlpShape* shape, const float* params, const float* x, float* y, int64_t n_rows,
                           void* stream);

/* ------------------------------------------------------------------------------------------------
 * Data-parallel gradient exchange over NVLink peer memory (one process per GPU, one node).
 * Replaces the two Model.reduce_parameters calls of DDPG._update (src/my_agents/ddpg.py:409-410 critic,
 * :431-432 actor; skrl: all_reduce(SUM) of the flat gradients, then / world_size) and the process-group
 * set-up of src/my_agents/base.py:88-92.  Every rank owns one exchange block in its HBM that holds its two
 * flat gradient buffers; the blocks are mapped into every peer with CUDA IPC and the Adam kernels of
 * ois_ddpg_update_phase read all ranks' gradients in place (csrc/comm.cuh states the protocol).  No host
 * involvement per update: the exchange is part of the captured update graph.
 * ---------------------------------------------------------------------------------------------- */
#define OIS_MAX_RANKS 8
#define OIS_IPC_HANDLE_BYTES 64
typedef struct OisComm OisComm;   /* opaque */

/* Allocate this rank's exchange block (zeroed): header + critic gradient + actor gradient (flat fp32 counts). */
int ois_comm_create(OisComm** out, int32_t rank, int32_t world, int64_t critic_params, int64_t actor_params);
/* The block's CUDA IPC handle (OIS_IPC_HANDLE_BYTES bytes) to hand to the other ranks by any host channel. */
int ois_comm_export(OisComm* comm, void* handle_out);
/* Map the peers' blocks: `handles` = world x OIS_IPC_HANDLE_BYTES bytes in rank order (own entry ignored). */
int ois_comm_connect(OisComm* comm, const void* handles, int32_t count);
/* Same-process variant: all ranks' OisComm objects live in this process (tests; several GPUs driven by one process). */
int ois_comm_connect_local(OisComm* comm, OisComm* const* all, int32_t count);
/* This rank's gradient buffer inside the block: net 0 = critic, 1 = actor.  These are the buffers to pass as
 * OisUpdateIO.critic_grad / actor_grad when OisUpdateIO.comm is set. */
void* ois_comm_grad(OisComm* comm, int32_t net);
/* words_host[0..1] = completed critic / actor exchanges, [2] = 0, or p + 1 when a barrier gave up waiting for rank p. */
int ois_comm_status(OisComm* comm, uint32_t* words_host);
/* Test hook: mark every rank as having published exchange `epoch` of `kind` in THIS rank's header (host write). */
int ois_comm_debug_publish(OisComm* comm, int32_t kind, uint32_t epoch);
int ois_comm_destroy(OisComm* comm);

typedef struct OisUpdateIO {
  /* sampled batch (ois_replay_gather outputs) */
  const float* states; const float* actions; const float* rewards; const float* next_states;
  const uint8_t* terminated; const uint8_t* truncated;
  /* parameters: flat fp32, see OisMlpShape */
  float* actor; float* actor_target; float* critic; float* critic_target;
  /* Adam state (torch.optim.Adam defaults), flat like the parameters */
  float* actor_m; float* actor_v; float* critic_m; float* critic_v;
  int64_t adam_step;          /* 1-based step count of THIS update */
  float actor_lr, critic_lr, beta1, beta2, adam_eps;
  float discount, polyak;
  /* gradient buffers (flat); with world_size > 1 the caller all-reduces them between the two phases */
  float* actor_grad; float* critic_grad;
  /* outputs for logging (ddpg.py:452-461): critic_loss, policy_loss, q max/min/mean, target max/min/mean */
  float* stats;               /* [8], nullable */
  const int64_t* adam_step_dev; /* optional device-resident 1-based step counter (CUDA-graph replay); overrides adam_step */
  OisComm* comm;              /* nullable; when set (world > 1): the Adam kernels of phases 1 and 2 first exchange the
                                 gradients of all ranks over peer memory (sum in rank order) — critic_grad / actor_grad must
                                 be ois_comm_grad(comm, 0 / 1) and grad_scale 1 / world */
  int32_t batch_rows;         /* rows of the batch tensors; 0 = the handle's batch_size.  A shorter batch (the replay
                                 memory holds fewer than batch_size transitions, which the reference trains on as is) is
                                 served by the fp32 path only */
} OisUpdateIO;

/* DDPG._update (ddpg.py:357-444) split at the two points where skrl all-reduces gradients:
 *   phase 0: scaler(train) on states & next_states, TD target, critic forward/backward -> critic_grad
 *   phase 1: Adam(critic, grad * grad_scale); actor forward, critic forward, policy backward -> actor_grad
 *   phase 2: Adam(actor, grad * grad_scale); polyak update of both targets
 * grad_scale = 1/world_size after an all-reduce(SUM) (skrl Model.reduce_parameters), else 1. */
int ois_ddpg_update_phase(OisAgent* agent, const OisScaler* scaler, const OisUpdateIO* io, int32_t phase,
                          float grad_scale, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Host in the loop: the trainer's loop body (src/my_trainers/base_KA_DDPG.py:175-219) is one captured CUDA graph per
 * policy step (act -> env step -> fused record_transition).  graph_exec is a cudaGraphExec_t.
 * ---------------------------------------------------------------------------------------------- */
int ois_graph_launch(void* graph_exec, void* stream);
/* Spin (no driver call) until the pinned host word has reached `expect` (wrap-safe >=), i.e. until the env kernel that
 * was given it as OisEnvIO.done_flag has completed that many steps.  timeout_ns <= 0: wait for ever. */
int ois_host_wait(const uint32_t* host_flag, uint32_t expect, int64_t timeout_ns);
/* ois_graph_launch followed by ois_host_wait in one call. */
int ois_graph_launch_wait(void* graph_exec, void* stream, const uint32_t* host_flag, uint32_t expect, int64_t timeout_ns);

/* ------------------------------------------------------------------------------------------------ */
const char* ois_last_error_string(void);
int ois_abi_version(void);
/* sizeof() of the ABI structs as this library was compiled, so a foreign-language mirror (ctypes, cgo ...)
 * can verify its layout: 0 OisEnvParams, 1 OisEnvState, 2 OisEnvIO, 3 OisMlpShape, 4 OisScaler, 5 OisActIO,
 * 6 OisUpdateIO, 7 OisStepCtl, 8 OisSampleCtl; -1 for an unknown index. */
int ois_struct_size(int which);
/* name of the SASS feature set compiled in, e.g. "sm_100a tcgen05+tma" */
const char* ois_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* OIS_B200_H */
