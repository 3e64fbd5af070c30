// Replay memory kernels: ring-slot append, gather by index, and a device-side
// replacement for the reference's host-side torch.randperm(len)[:batch].
//
// Memory layout follows skrl Memory: every tensor is (memory_size, num_envs, dim), sampled through its
// flattened (memory_size*num_envs, dim) view (train/train_kaddpg_1d_nofifo_random_RLIL.py:130-160).
#include "ois_common.cuh"

namespace ois {

constexpr int kThreads = 256;

// one launch instead of six copy_ calls: thread t copies float #t of every tensor's slot row
__global__ void __launch_bounds__(kThreads)
replay_append_kernel(float* __restrict__ ms, float* __restrict__ ma, float* __restrict__ mr, float* __restrict__ mn,
                     uint8_t* __restrict__ mt, uint8_t* __restrict__ mu, const float* __restrict__ s,
                     const float* __restrict__ a, const float* __restrict__ r, const float* __restrict__ ns,
                     const uint8_t* __restrict__ te, const uint8_t* __restrict__ tr, int64_t n, int obs_dim,
                     int act_dim) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t t = t0; t < n * obs_dim; t += stride) { ms[t] = s[t]; mn[t] = ns[t]; }
  for (int64_t t = t0; t < n * act_dim; t += stride) ma[t] = a[t];
  for (int64_t t = t0; t < n; t += stride) { mr[t] = r[t]; mt[t] = te[t]; mu[t] = tr ? tr[t] : 0; }
}

// One warp-lane group per sampled row.  Rows are 24/4..16/4/1/1 bytes: sector-granular random reads, so the
// gather is bound by L2/HBM sector throughput (32 B), not by bytes.  16 lanes cover one row of every tensor.
__global__ void __launch_bounds__(kThreads)
replay_gather_kernel(const float* __restrict__ ms, const float* __restrict__ ma, const float* __restrict__ mr,
                     const float* __restrict__ mn, const uint8_t* __restrict__ mt, const uint8_t* __restrict__ mu,
                     const int64_t* __restrict__ idx, int64_t batch, int64_t rows, int obs_dim, int act_dim,
                     float* __restrict__ os, float* __restrict__ oa, float* __restrict__ orw, float* __restrict__ on,
                     uint8_t* __restrict__ ot, uint8_t* __restrict__ ou) {
  const int lane = threadIdx.x & 15;
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) >> 4;
  for (int64_t b = group; b < batch; b += ngroups) {
    int64_t j = idx[b];
    if (j < 0 || j >= rows) j = 0;   // defensive: indices come from the caller
    for (int k = lane; k < obs_dim; k += 16) {
      os[b * obs_dim + k] = ms[j * obs_dim + k];
      on[b * obs_dim + k] = mn[j * obs_dim + k];
    }
    for (int k = lane; k < act_dim; k += 16) oa[b * act_dim + k] = ma[j * act_dim + k];
    if (lane == 15) {
      orw[b] = mr[j];
      ot[b] = mt[j];
      ou[b] = mu[j];
    }
  }
}

// `batch` distinct indices in [0, len): positions 0..batch-1 of a keyed pseudo-random permutation of
// [0, 2^k) (4-round Feistel on k bits, k even, 2^k >= len), cycle-walked back into [0, len).
// Distinctness is structural (a permutation), like randperm(len)[:batch]; the distribution is that of the
// Feistel network, not of Fisher-Yates: statistical tests only, parity tests supply indices.
__device__ __forceinline__ uint64_t feistel(uint64_t x, int half_bits, const uint32_t key[4]) {
  const uint64_t mask = (1ull << half_bits) - 1;
  uint64_t l = x >> half_bits, r = x & mask;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    uint64_t f = (r + key[i]) * 0x9E3779B97F4A7C15ull;
    f ^= f >> 29;
    f *= 0xBF58476D1CE4E5B9ull;
    f ^= f >> 32;
    const uint64_t nl = r;
    r = l ^ (f & mask);
    l = nl;
  }
  return (l << half_bits) | r;
}

__global__ void replay_sample_kernel(int64_t* __restrict__ idx, int64_t batch, int64_t len, uint64_t seed,
                                     uint64_t offset) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  int bits = 2;
  while ((1ull << bits) < (uint64_t)len) bits += 2;
  const uint4 k4 = Philox(seed)(offset, 0x5eedull);
  const uint32_t key[4] = {k4.x, k4.y, k4.z, k4.w};
  uint64_t x = (uint64_t)b;
  do { x = feistel(x, bits / 2, key); } while (x >= (uint64_t)len);
  idx[b] = (int64_t)x;
}

// Sampler and gather in ONE launch, with everything that changes from one update to the next read from (and advanced in)
// a device-resident OisSampleCtl — so a captured CUDA graph of DDPG._update needs no launch outside the graph.  The
// indices are the ones replay_sample_kernel draws for (len, seed, draws); they are also stored to idx (Memory's
// sampling_indexes).  The last CTA to finish moves the block to the next update.
__global__ void __launch_bounds__(kThreads)
replay_sample_gather_kernel(const float* __restrict__ ms, const float* __restrict__ ma, const float* __restrict__ mr,
                            const float* __restrict__ mn, const uint8_t* __restrict__ mt, const uint8_t* __restrict__ mu,
                            int64_t* __restrict__ idx, int64_t batch, int obs_dim, int act_dim, float* __restrict__ os,
                            float* __restrict__ oa, float* __restrict__ orw, float* __restrict__ on, uint8_t* __restrict__ ot,
                            uint8_t* __restrict__ ou, uint64_t seed, OisSampleCtl* __restrict__ ctl) {
  pdl_wait();
  if (threadIdx.x == 0) pdl_trigger();
  const int64_t len = ctl->len;
  const uint64_t draws = ctl->draws;
  int bits = 2;
  while ((1ull << bits) < (uint64_t)len) bits += 2;
  const uint4 k4 = Philox(seed)(draws, 0x5eedull);
  const uint32_t key[4] = {k4.x, k4.y, k4.z, k4.w};
  const int lane = threadIdx.x & 15;
  const int64_t group = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 4;
  const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) >> 4;
  for (int64_t b = group; b < batch; b += ngroups) {
    uint64_t x = (uint64_t)b;
    do { x = feistel(x, bits / 2, key); } while (x >= (uint64_t)len);
    const int64_t j = (int64_t)x;
    if (lane == 0) idx[b] = j;
    for (int k = lane; k < obs_dim; k += 16) {
      os[b * obs_dim + k] = ms[j * obs_dim + k];
      on[b * obs_dim + k] = mn[j * obs_dim + k];
    }
    for (int k = lane; k < act_dim; k += 16) oa[b * act_dim + k] = ma[j * act_dim + k];
    if (lane == 15) {
      orw[b] = mr[j];
      ot[b] = mt[j];
      ou[b] = mu[j];
    }
  }
  __syncthreads();      // every thread of this CTA has read the block
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ctl->ticket, 1u) == gridDim.x - 1) {
      ctl->ticket = 0;
      ctl->draws = draws + 1;
      ctl->adam_step += 1;
    }
  }
}

}  // namespace ois
using namespace ois;

namespace ois {
__global__ void sample_ctl_set_kernel(OisSampleCtl* ctl, int64_t len, uint64_t draws, int64_t adam_step) {
  ctl->len = len;
  ctl->draws = draws;
  ctl->adam_step = adam_step;
  ctl->ticket = 0;
}
}  // namespace ois

// Values travel as kernel arguments (captured at launch), so the host may call this again before the first call has
// executed; a plain stream-ordered predecessor of the next ois_replay_sample_gather.
extern "C" int ois_sample_ctl_set(OisSampleCtl* ctl, int64_t len, uint64_t draws, int64_t adam_step, void* stream) {
  OIS_CHECK_ARG(ctl && len >= 0 && adam_step >= 0, "bad arguments");
  sample_ctl_set_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(ctl, len, draws, adam_step);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

extern "C" int ois_replay_sample_gather(const float* mem_states, const float* mem_actions, const float* mem_rewards,
                                        const float* mem_next_states, const uint8_t* mem_terminated,
                                        const uint8_t* mem_truncated, int64_t* idx, int64_t batch, int32_t obs_dim,
                                        int32_t act_dim, float* out_states, float* out_actions, float* out_rewards,
                                        float* out_next_states, uint8_t* out_terminated, uint8_t* out_truncated,
                                        uint64_t philox_seed, OisSampleCtl* ctl, void* stream) {
  OIS_CHECK_ARG(mem_states && mem_actions && mem_rewards && mem_next_states && mem_terminated && mem_truncated,
                "null memory tensor");
  OIS_CHECK_ARG(out_states && out_actions && out_rewards && out_next_states && out_terminated && out_truncated && idx && ctl,
                "null output tensor / control block");
  OIS_CHECK_ARG(batch > 0 && obs_dim > 0 && act_dim > 0, "bad sizes");
  int64_t blocks = (batch * 16 + kThreads - 1) / kThreads;
  if (blocks > (int64_t)num_sms() * 8) blocks = (int64_t)num_sms() * 8;
  OIS_CHECK_CUDA(launch_pdl(replay_sample_gather_kernel, (int)blocks, kThreads, 0, (cudaStream_t)stream, mem_states, mem_actions,
                            mem_rewards, mem_next_states, mem_terminated, mem_truncated, idx, batch, (int)obs_dim, (int)act_dim,
                            out_states, out_actions, out_rewards, out_next_states, out_terminated, out_truncated, philox_seed, ctl));
  return OIS_OK;
}

extern "C" int ois_replay_append(float* mem_states, float* mem_actions, float* mem_rewards, float* mem_next_states,
                                 uint8_t* mem_terminated, uint8_t* mem_truncated, int64_t memory_index, int64_t n_envs,
                                 int32_t obs_dim, int32_t act_dim, const float* states, const float* actions,
                                 const float* rewards, const float* next_states, const uint8_t* terminated,
                                 const uint8_t* truncated, void* stream) {
  OIS_CHECK_ARG(mem_states && mem_actions && mem_rewards && mem_next_states && mem_terminated && mem_truncated,
                "null memory tensor");
  OIS_CHECK_ARG(states && actions && rewards && next_states && terminated, "null sample tensor");
  OIS_CHECK_ARG(memory_index >= 0 && n_envs >= 0 && obs_dim > 0 && act_dim > 0, "bad sizes");
  if (n_envs == 0) return OIS_OK;
  const int64_t row = memory_index * n_envs;
  int64_t blocks = (n_envs * obs_dim + kThreads - 1) / kThreads;
  if (blocks > (int64_t)num_sms() * 8) blocks = (int64_t)num_sms() * 8;
  replay_append_kernel<<<(int)blocks, kThreads, 0, (cudaStream_t)stream>>>(
      mem_states + row * obs_dim, mem_actions + row * act_dim, mem_rewards + row, mem_next_states + row * obs_dim,
      mem_terminated + row, mem_truncated + row, states, actions, rewards, next_states, terminated, truncated, n_envs,
      obs_dim, act_dim);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

extern "C" int ois_replay_gather(const float* mem_states, const float* mem_actions, const float* mem_rewards,
                                 const float* mem_next_states, const uint8_t* mem_terminated,
                                 const uint8_t* mem_truncated, const int64_t* idx, int64_t batch, int64_t mem_rows,
                                 int32_t obs_dim, int32_t act_dim, float* out_states, float* out_actions,
                                 float* out_rewards, float* out_next_states, uint8_t* out_terminated,
                                 uint8_t* out_truncated, void* stream) {
  OIS_CHECK_ARG(mem_states && mem_actions && mem_rewards && mem_next_states && mem_terminated && mem_truncated,
                "null memory tensor");
  OIS_CHECK_ARG(out_states && out_actions && out_rewards && out_next_states && out_terminated && out_truncated && idx,
                "null output tensor");
  OIS_CHECK_ARG(batch >= 0 && mem_rows > 0 && obs_dim > 0 && act_dim > 0, "bad sizes");
  if (batch == 0) return OIS_OK;
  int64_t blocks = (batch * 16 + kThreads - 1) / kThreads;
  if (blocks > (int64_t)num_sms() * 8) blocks = (int64_t)num_sms() * 8;
  replay_gather_kernel<<<(int)blocks, kThreads, 0, (cudaStream_t)stream>>>(
      mem_states, mem_actions, mem_rewards, mem_next_states, mem_terminated, mem_truncated, idx, batch, mem_rows,
      obs_dim, act_dim, out_states, out_actions, out_rewards, out_next_states, out_terminated, out_truncated);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}

extern "C" int ois_replay_sample_indices(int64_t* idx, int64_t batch, int64_t len, uint64_t philox_seed,
                                         uint64_t philox_offset, void* stream) {
  OIS_CHECK_ARG(idx && batch >= 0 && len > 0, "bad arguments");
  OIS_CHECK_ARG(batch <= len, "batch larger than the memory length (sampling is without replacement)");
  if (batch == 0) return OIS_OK;
  replay_sample_kernel<<<(int)((batch + 127) / 128), 128, 0, (cudaStream_t)stream>>>(idx, batch, len, philox_seed,
                                                                                    philox_offset);
  OIS_CHECK_LAUNCH();
  return OIS_OK;
}
