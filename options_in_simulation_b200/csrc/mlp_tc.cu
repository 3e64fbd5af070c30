// bf16 tcgen05 + TMA fused actor MLP (placeholder until the tensor-core kernel lands: fails loudly).
#include "agent.cuh"
namespace ois {
int tc_agent_init(OisAgent*, cudaStream_t) { set_error("bf16 tcgen05 path not built in this revision"); return OIS_ERR_UNSUPPORTED; }
void tc_agent_free(OisAgent*) {}
void tc_invalidate_weights(OisAgent*) {}
int tc_actor_forward(OisAgent*, const float*, const OisScaler*, const float*, float*, int64_t, cudaStream_t) {
  set_error("bf16 tcgen05 path not built in this revision");
  return OIS_ERR_UNSUPPORTED;
}
}  // namespace ois
