// Host side of the peer-memory gradient exchange (comm.cuh states the protocol): allocation of this rank's exchange
// block, CUDA IPC export / import of the blocks of all ranks of one node, and the tables the exchanging Adam kernels take.
// Reference: skrl's process-group set-up in src/my_agents/base.py:88-92 and the two gradient reductions of
// src/my_agents/ddpg.py:409-410,431-432.  Handles travel between the processes as opaque 64-byte strings; how they
// travel (torch.distributed all_gather_object, a file, MPI ...) is the caller's business.
#include <new>
#include <string.h>
#include "comm.cuh"

static_assert(sizeof(cudaIpcMemHandle_t) == OIS_IPC_HANDLE_BYTES, "OIS_IPC_HANDLE_BYTES must match cudaIpcMemHandle_t");

namespace ois {

CommArgs comm_args(const OisComm* c, int kind) {
  CommArgs a = {};
  if (!c) return a;
  a.world = c->world;
  a.rank = c->rank;
  a.kind = kind;
  a.header = (uint32_t*)c->block;
  for (int p = 0; p < c->world; ++p) {
    a.peer_flag[p] = (uint32_t*)c->peer[p] + kCommFlags + kind * OIS_MAX_RANKS + c->rank;
    a.peer_grad[p] = (const float*)((const char*)c->peer[p] + c->off_grad[kind]);
  }
  return a;
}

}  // namespace ois

using namespace ois;

extern "C" int ois_comm_create(OisComm** out, int32_t rank, int32_t world, int64_t critic_params, int64_t actor_params) {
  OIS_CHECK_ARG(out, "null argument");
  OIS_CHECK_ARG(world >= 1 && world <= OIS_MAX_RANKS && rank >= 0 && rank < world, "need 0 <= rank < world <= OIS_MAX_RANKS");
  OIS_CHECK_ARG(critic_params > 0 && actor_params > 0, "parameter counts must be positive");
  OisComm* c = new (std::nothrow) OisComm();
  OIS_CHECK_ARG(c, "out of host memory");
  c->rank = rank;
  c->world = world;
  c->n[0] = critic_params;
  c->n[1] = actor_params;
  auto up = [](size_t b) { return (b + 255) / 256 * 256; };
  c->off_grad[0] = OIS_COMM_HEADER_BYTES;
  c->off_grad[1] = c->off_grad[0] + up((size_t)critic_params * sizeof(float));
  c->block_bytes = c->off_grad[1] + up((size_t)actor_params * sizeof(float));
  cudaError_t e = cudaMalloc(&c->block, c->block_bytes);
  if (e == cudaSuccess) e = cudaMemset(c->block, 0, c->block_bytes);
  if (e != cudaSuccess) {
    const size_t bytes = c->block_bytes;
    if (c->block) cudaFree(c->block);
    delete c;
    set_error("ois_comm_create: cudaMalloc/cudaMemset(%zu) -> %s", bytes, cudaGetErrorString(e));
    return OIS_ERR_CUDA;
  }
  c->peer[rank] = c->block;
  c->connected = world == 1;
  *out = c;
  return OIS_OK;
}

extern "C" int ois_comm_export(OisComm* c, void* handle_out) {
  OIS_CHECK_ARG(c && handle_out, "null argument");
  cudaIpcMemHandle_t h;
  OIS_CHECK_CUDA(cudaIpcGetMemHandle(&h, c->block));
  memcpy(handle_out, &h, sizeof(h));
  return OIS_OK;
}

extern "C" int ois_comm_connect(OisComm* c, const void* handles, int32_t count) {
  OIS_CHECK_ARG(c && handles, "null argument");
  OIS_CHECK_ARG(count == c->world, "need one handle per rank, in rank order");
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank || c->peer[p]) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)handles + (size_t)p * sizeof(h), sizeof(h));
    void* ptr = nullptr;
    OIS_CHECK_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    c->peer[p] = ptr;
    c->opened[p] = true;
  }
  c->connected = true;
  return OIS_OK;
}

// Same-process stand-in for ois_comm_connect: the blocks of all ranks are plain device pointers the caller owns (several
// OisComm objects in one process, on one GPU or on peer-enabled GPUs).  Tests use it to check the reduction arithmetic
// on a single-GPU box.
extern "C" int ois_comm_connect_local(OisComm* c, OisComm* const* all, int32_t count) {
  OIS_CHECK_ARG(c && all, "null argument");
  OIS_CHECK_ARG(count == c->world, "need one comm per rank, in rank order");
  for (int p = 0; p < c->world; ++p) {
    OIS_CHECK_ARG(all[p] && all[p]->world == c->world && all[p]->rank == p && all[p]->n[0] == c->n[0] && all[p]->n[1] == c->n[1],
                  "comm objects do not describe the same world");
    c->peer[p] = all[p]->block;
  }
  c->connected = true;
  return OIS_OK;
}

extern "C" void* ois_comm_grad(OisComm* c, int32_t net) {
  if (!c || net < 0 || net > 1) return nullptr;
  return (char*)c->block + c->off_grad[net];
}

// Host-side write of flags[kind][*] := epoch in THIS rank's header, as if every rank had published (tests on a
// single-GPU box, where kernels that wait for one another must not be run as separate launches).
extern "C" int ois_comm_debug_publish(OisComm* c, int32_t kind, uint32_t epoch) {
  OIS_CHECK_ARG(c && kind >= 0 && kind <= 1, "bad argument");
  uint32_t v[OIS_MAX_RANKS];
  for (int p = 0; p < OIS_MAX_RANKS; ++p) v[p] = epoch;
  OIS_CHECK_CUDA(cudaMemcpy((uint32_t*)c->block + kCommFlags + kind * OIS_MAX_RANKS, v, sizeof(v), cudaMemcpyHostToDevice));
  return OIS_OK;
}

// words[0..1] = completed exchanges (critic, actor), words[2] = status (0 ok; p + 1: the barrier gave up on rank p)
extern "C" int ois_comm_status(OisComm* c, uint32_t* words_host) {
  OIS_CHECK_ARG(c && words_host, "null argument");
  uint32_t h[40];
  OIS_CHECK_CUDA(cudaMemcpy(h, c->block, sizeof(h), cudaMemcpyDeviceToHost));
  words_host[0] = h[kCommEpoch];
  words_host[1] = h[kCommEpoch + 1];
  words_host[2] = h[kCommStatus];
  return OIS_OK;
}

extern "C" int ois_comm_destroy(OisComm* c) {
  if (!c) return OIS_OK;
  for (int p = 0; p < c->world; ++p)
    if (c->opened[p] && c->peer[p]) cudaIpcCloseMemHandle(c->peer[p]);
  if (c->block) cudaFree(c->block);
  delete c;
  return OIS_OK;
}
