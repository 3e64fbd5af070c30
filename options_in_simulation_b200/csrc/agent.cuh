// Internal agent handle shared by agent.cu (fp32 path, update) and mlp_tc.cu (bf16 tcgen05 path).
#pragma once
#include "ois_common.cuh"

struct OisAgent {
  OisMlpShape actor, critic;
  int64_t max_rows = 0;
  int batch = 0;
  int precision = 0;
  void* ws = nullptr;
  size_t ws_bytes = 0;
  // update workspace (batch rows)
  float *u_sn = nullptr, *u_s2n = nullptr, *u_x = nullptr, *u_a = nullptr, *u_q = nullptr, *u_q2 = nullptr;
  float *u_y = nullptr, *u_dq = nullptr, *u_dx = nullptr, *u_dz = nullptr;
  float *ah0 = nullptr, *ah1 = nullptr, *ah2 = nullptr, *ch0 = nullptr, *ch1 = nullptr, *ch2 = nullptr;
  float *d0 = nullptr, *d1 = nullptr, *d2 = nullptr;
  // act workspace
  float *xn = nullptr, *h0 = nullptr, *h1 = nullptr, *h2 = nullptr;
  // bf16 tensor-core path (mlp_tc.cu)
  void* tc = nullptr;
  // bf16 tensor-core update path (update_tc.cu)
  void* tcu = nullptr;
};

namespace ois {

struct MlpOffsets {
  int64_t w[4], b[4], total;
  int in[4], out[4];
};
MlpOffsets mlp_offsets(const OisMlpShape& s);

int mlp_forward_fp32(const OisMlpShape& sh, const float* P, const float* x, int ldx, float* h0, float* h1, float* h2,
                     float* y, int64_t rows, cudaStream_t s);

// bf16 tcgen05 + TMA fused actor (mlp_tc.cu)
int tc_agent_init(OisAgent* a, cudaStream_t s);
void tc_agent_free(OisAgent* a);
void tc_invalidate_weights(OisAgent* a);
int tc_sync_image(OisAgent* a, const float* actor_params, int64_t rows, cudaStream_t s);
// the small-tile transposed kernel (mlp_tct.cu): 64 envs per CTA, weights from the update's tile-blocked image
int tct_actor_forward(OisAgent* a, const uint8_t* image, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                      int64_t rows, cudaStream_t s);
// scaler -> 4-layer MLP -> (OU noise, scale, clamp when io->ou_state) in ONE launch
int tc_actor_forward(OisAgent* a, const float* actor_params, const OisScaler* scaler, const OisActIO* io, int64_t rows,
                     cudaStream_t s);

// the fused policy step (mlp_tct.cu): the transposed actor kernel with the env step of each tile run by the CTA that
// produced its actions; rows must satisfy tct_can_fuse_env
bool tc_use_transposed(const OisAgent* a, int64_t rows);
bool tct_can_fuse_env(int64_t rows);
int tct_rollout_forward(OisAgent* a, const uint8_t* image, const float* actor_params, const OisScaler* scaler, const OisActIO* io,
                        const OisEnvParams* ep, const OisEnvState* es, const OisEnvIO* eio, int64_t rows, cudaStream_t s);
// Called by ois_env_step (env_step.cu): if ois_agent_act deferred an actor launch on this thread (between
// ois_rollout_fuse_begin / _end) and this env step consumes exactly its actions, both run as ONE launch (returns 1);
// otherwise the deferred actor launch is issued first and the caller launches the env kernel itself (returns 0).
int rollout_try_fuse(const OisEnvParams* params, const OisEnvState* state, const OisEnvIO* io, int64_t n, cudaStream_t s);

}  // namespace ois
