// bf16 tcgen05 building blocks of DDPG._update (update_tc.cu), called from ois_ddpg_update_phase (agent.cu).
#pragma once
#include "agent.cuh"

namespace ois {

// byte offsets of the four tile-blocked bf16 weight matrices inside one network's image
constexpr size_t kImgW1 = 0;                                  // [512, 16]   row block 128
constexpr size_t kImgW2 = kImgW1 + 512 * 16 * 2;              // [512, 512]  row block 128
constexpr size_t kImgW3 = kImgW2 + 512 * 512 * 2;             // [128, 512]  row block 128
constexpr size_t kImgW4 = kImgW3 + 128 * 512 * 2;             // [16, 128]   row block 16
constexpr size_t kImgBytes = kImgW4 + 16 * 128 * 2;

enum { TCU_ACTOR = 0, TCU_ACTOR_T = 1, TCU_CRITIC = 2, TCU_CRITIC_T = 3 };        // weight images
enum { TCU_ACTS_ACTOR = 0, TCU_ACTS_CRITIC = 1, TCU_ACTS_SCRATCH = 2, TCU_DZ = 3 };  // activation sets / pack targets

int tcu_init(OisAgent* a);
void tcu_free(OisAgent* a);
void tcu_invalidate(OisAgent* a, int which);   // which < 0: all four images
// fp32 master -> bf16 image of network `net` (no-op while the image is valid for this parameter buffer)
int tcu_image(OisAgent* a, int net, const float* P32, cudaStream_t s);
// [B, w1] (columns col1.. of s1) ++ [B, w2] -> the bf16 input of activation set `dst` (or the dz buffer for TCU_DZ);
// with act_out the first block is multiplied by the output activation's derivative at act_out
int tcu_pack_x(OisAgent* a, int dst, const float* s1, int ld1, int col1, int w1, const float* s2, int ld2, int w2,
               const float* act_out, int act, cudaStream_t s);
// y_i = mlp_i(x_i) for up to two independent networks, layer by layer in grouped launches; hidden activations of set
// acts[i] are kept (bf16) for the backward
int tcu_forward(OisAgent* a, int count, const int* nets, const float* const* P32, const int* acts, float* const* y,
                cudaStream_t s);
// backward of network `net` from the dz buffer through activation set `acts`; G: flat fp32 parameter gradients
// (nullable), dx: [B, in_dim] fp32 input gradient (nullable)
int tcu_backward(OisAgent* a, int net, float* G, int acts, float* dx, cudaStream_t s);

}  // namespace ois
