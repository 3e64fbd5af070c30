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

// byte offset of chunk (r, c8) of a tile-blocked bf16 matrix with C8 column chunks and RB rows per block
__host__ __device__ inline size_t tcu_chunk_off(int r, int c8, int C8, int rb) {
  return ((size_t)((r / rb) * C8 + c8) * rb + (r % rb)) * 16;
}
// byte offset inside a network's image of flat parameter i (nn.Sequential order, 512-512-128 hidden); -1 for biases
__host__ __device__ inline int64_t tcu_img_elem_off(int64_t i, int in_dim, int out_dim) {
  const int64_t b0 = (int64_t)512 * in_dim, w1 = b0 + 512, b1 = w1 + 512 * 512, w2 = b1 + 512, b2 = w2 + 128 * 512;
  const int64_t w3 = b2 + 128, b3 = w3 + (int64_t)out_dim * 128;
  if (i < b0) { const int r = (int)(i / in_dim), c = (int)(i % in_dim); return (int64_t)(kImgW1 + tcu_chunk_off(r, c >> 3, 2, 128)) + (c & 7) * 2; }
  if (i < w1) return -1;
  if (i < b1) { const int j = (int)(i - w1), r = j >> 9, c = j & 511; return (int64_t)(kImgW2 + tcu_chunk_off(r, c >> 3, 64, 128)) + (c & 7) * 2; }
  if (i < w2) return -1;
  if (i < b2) { const int j = (int)(i - w2), r = j >> 9, c = j & 511; return (int64_t)(kImgW3 + tcu_chunk_off(r, c >> 3, 64, 128)) + (c & 7) * 2; }
  if (i < w3) return -1;
  if (i < b3) { const int j = (int)(i - w3), r = j >> 7, c = j & 127; return (int64_t)(kImgW4 + tcu_chunk_off(r, c >> 3, 16, 16)) + (c & 7) * 2; }
  return -1;
}
#ifdef __CUDACC__
__device__ inline uint16_t tcu_bf16_bits(float v) {
  uint16_t r;
  asm("cvt.rn.bf16.f32 %0, %1;" : "=h"(r) : "f"(v));
  return r;
}
__device__ inline void tcu_img_store(uint8_t* img, int64_t i, float v, int in_dim, int out_dim) {
  const int64_t off = tcu_img_elem_off(i, in_dim, out_dim);
  if (off >= 0) *reinterpret_cast<uint16_t*>(img + off) = tcu_bf16_bits(v);
}
#endif

enum { TCU_ACTOR = 0, TCU_ACTOR_T = 1, TCU_CRITIC = 2, TCU_CRITIC_T = 3 };        // weight images
enum { TCU_ACTS_ACTOR = 0, TCU_ACTS_CRITIC = 1, TCU_ACTS_SCRATCH = 2, TCU_ACTS_POLICY = 3, TCU_DZ = 4 };  // activation sets / pack targets

int tcu_init(OisAgent* a);
void tcu_free(OisAgent* a);
void tcu_invalidate(OisAgent* a, int which);   // which < 0: all four images
// fp32 master -> bf16 image of network `net` (no-op while the image is valid for this parameter buffer)
int tcu_image(OisAgent* a, int net, const float* P32, cudaStream_t s);
// the image of `net` if it is complete for parameter buffer P32 (then a fused Adam / polyak kernel may keep it current
// element by element, tcu_img_store), else null
uint8_t* tcu_image_if_valid(OisAgent* a, int net, const float* P32);
// the bf16 input buffer of activation set `acts` (tile-blocked [B, 16]) / the dz buffer for TCU_DZ
uint8_t* tcu_x_buffer(OisAgent* a, int which);
// [B, w1] (columns col1.. of s1) ++ [B, w2] -> the bf16 input of activation set `dst` (or the dz buffer for TCU_DZ);
// with act_out the first block is multiplied by the output activation's derivative at act_out; with q the same launch
// also writes the policy-loss statistics of ddpg.py:423 (stats[1..4] = loss, max q, min q, mean q)
int tcu_pack_x(OisAgent* a, int dst, const float* s1, int ld1, int col1, int w1, const float* s2, int ld2, int w2,
               const float* act_out, int act, const float* q, float* stats, cudaStream_t s);
// y_i = mlp_i(x_i) for up to two independent networks, layer by layer in grouped launches; hidden activations of set
// acts[i] are kept (bf16) for the backward
// x_out[i] (nullable): the outputs are ALSO written, as bf16, into columns x_col.. of that tile-blocked [B, 16] buffer
// (the actor's action lands directly in the critic's input)
int tcu_forward(OisAgent* a, int count, const int* nets, const float* const* P32, const int* acts, float* const* y,
                uint8_t* const* x_out, int x_col, cudaStream_t s);
// backward of network `net` from the dz buffer through activation set `acts`; G: flat fp32 parameter gradients
// (nullable), dx: [B, in_dim] fp32 input gradient (nullable); policy_dz: start from the constant -1/B gradient of the
// policy loss instead of the dz buffer
// have_d2: the layer-3 gradient d2 is already in tcu_d2_buffer (written together with dz by the fused loss kernels of
// agent.cu), so the K = 16 data-gradient launch is skipped
int tcu_backward(OisAgent* a, int net, float* G, int acts, float* dx, bool policy_dz, bool have_d2, cudaStream_t s);
// the policy loss's trip through the critic (ddpg.py:420-429): q = critic(x of set `acts`) and dx = d(-mean q)/dx
int tcu_policy_pass(OisAgent* a, int net, const float* P32, int acts, float* q, float* dx, cudaStream_t s);
uint8_t* tcu_d2_buffer(OisAgent* a);
const uint8_t* tcu_h2_buffer(OisAgent* a, int acts);
const uint8_t* tcu_w4_image(OisAgent* a, int net);

}  // namespace ois
