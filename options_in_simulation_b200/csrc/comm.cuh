// Data-parallel gradient exchange over NVLink peer memory (reference: the two
// Model.reduce_parameters calls of src/my_agents/ddpg.py:409-410 and :431-432, which skrl implements as one
// NCCL all-reduce(SUM) of the concatenated gradients followed by a division by the world size).
//
// One process per GPU.  Every rank owns ONE exchange block in its own HBM (cudaMalloc, exported with CUDA IPC and mapped
// by all peers over NVLink 5 / NVSwitch):
//
//     [ header: flags[2][8] written by the peers | local words ]  [ critic gradient, flat fp32 ]  [ actor gradient ]
//
// The backward kernels of DDPG._update write this rank's gradients straight into the block (it IS the gradient buffer);
// the Adam kernel that consumes them is the collective:
//   1. the first CTA of the Adam launch to start publishes "my gradient k is complete" by storing the exchange's epoch
//      into flags[k][rank] of EVERY rank's block (release at system scope, after this GPU's backward kernels completed);
//   2. every CTA waits until all `world` entries of its own flags[k][*] have reached the epoch (acquire, system scope);
//   3. every thread sums element i of all ranks' gradient buffers in RANK ORDER (remote loads over NVLink; the same
//      order on every rank, hence bit-identical sums and bit-identical weights everywhere), scales by 1/world and
//      applies the Adam step — no reduced copy is ever written;
//   4. the last CTA to finish advances the epoch.
// There is exactly one barrier per exchange: a rank overwrites gradient k for update u+1 only after it has passed the
// barrier of the OTHER exchange of update u, which every peer reaches only after it has finished reading gradient k of
// update u (stream order), so the two exchanges of an update fence each other's buffers.
// Nothing here needs the host: the kernels are captured into the update's CUDA graph like any other launch.
#pragma once
#include "ois_common.cuh"

#define OIS_COMM_HEADER_BYTES 1024

struct OisComm {
  int rank = 0, world = 1;
  void* block = nullptr;                 // this rank's exchange block
  size_t block_bytes = 0;
  size_t off_grad[2] = {0, 0};           // byte offsets of the critic / actor gradient inside a block
  int64_t n[2] = {0, 0};                 // parameter counts
  void* peer[OIS_MAX_RANKS] = {};        // device-visible base of every rank's block (peer[rank] == block)
  bool opened[OIS_MAX_RANKS] = {};       // mapped with cudaIpcOpenMemHandle (to be closed)
  bool connected = false;
};

namespace ois {

// header words (uint32) inside a block
constexpr int kCommFlags = 0;            // [2][OIS_MAX_RANKS]  peer-written: epoch of rank r's gradient k
constexpr int kCommEpoch = 32;           // [2] local: completed exchanges of kind k
constexpr int kCommArrive = 34;          // [2] local: CTAs of the running Adam launch that have started
constexpr int kCommDepart = 36;          // [2] local: ... that have finished
constexpr int kCommStatus = 38;          // local: != 0 after a barrier timed out (a peer never arrived)

// by-value kernel argument of the exchanging Adam kernels; world == 0: no exchange (single process)
struct CommArgs {
  int world, rank, kind;
  uint32_t* header;                       // this rank's header
  uint32_t* peer_flag[OIS_MAX_RANKS];     // &flags[kind][rank] inside rank p's header
  const float* peer_grad[OIS_MAX_RANKS];  // gradient `kind` of rank p
};

CommArgs comm_args(const OisComm* c, int kind);

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// element i of rank p's gradient; system-scope relaxed (never served from this SM's L1)
__device__ __forceinline__ float ld_peer(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long comm_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// Steps 1 and 2 of the exchange; every thread of every CTA calls it (after pdl_wait).  Returns with all ranks' gradients
// of kind c.kind complete and visible.
__device__ __forceinline__ void comm_barrier(const CommArgs& c) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) {
    s_epoch = c.header[kCommEpoch + c.kind] + 1u;
    // first CTA of this launch to get here publishes; which CTA that is does not depend on the block scheduler's order
    if (atomicAdd(&c.header[kCommArrive + c.kind], 1u) == 0u) s_epoch |= 0x80000000u;
  }
  __syncthreads();
  const bool publish = (s_epoch & 0x80000000u) != 0u;
  const uint32_t epoch = s_epoch & 0x7FFFFFFFu;
  if (publish && (int)threadIdx.x < c.world) {
    __threadfence_system();                        // this GPU's gradient writes (earlier kernels) before the flag
    st_release_sys(c.peer_flag[threadIdx.x], epoch);
  }
  if ((int)threadIdx.x < c.world) {
    const uint32_t* f = c.header + kCommFlags + c.kind * OIS_MAX_RANKS + threadIdx.x;
    const unsigned long long t0 = comm_now_ns();
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
      if (comm_now_ns() - t0 > 20000000000ull) {   // 20 s: a peer died; flag it and fall through (no hang)
        atomicExch(&c.header[kCommStatus], 1u + (uint32_t)threadIdx.x);
        break;
      }
    }
  }
  __syncthreads();
}

// Step 4; every CTA calls it once after its last read of a peer buffer.
__device__ __forceinline__ void comm_finish(const CommArgs& c) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&c.header[kCommDepart + c.kind], 1u) == gridDim.x - 1u) {
      c.header[kCommArrive + c.kind] = 0u;
      c.header[kCommDepart + c.kind] = 0u;
      c.header[kCommEpoch + c.kind] += 1u;
      __threadfence();
    }
  }
}

// Step 3 for one element: sum over the ranks in rank order.
__device__ __forceinline__ float comm_reduced(const CommArgs& c, int64_t i) {
  float v[OIS_MAX_RANKS];
#pragma unroll
  for (int p = 0; p < OIS_MAX_RANKS; ++p)
    if (p < c.world) v[p] = ld_peer(c.peer_grad[p] + i);     // all loads in flight before the first add
  float s = v[0];
#pragma unroll
  for (int p = 1; p < OIS_MAX_RANKS; ++p)
    if (p < c.world) s = __fadd_rn(s, v[p]);
  return s;
}
#endif

}  // namespace ois
