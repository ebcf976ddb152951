// chain_tc.cu — tcgen05 / TMEM fused per-point MLP chains (bf16 operands, fp32 accumulate).
#include "model.cuh"
namespace pcp {
int model_forward_bf16(pcp_model *, const float *, int, int64_t, uint32_t, float *, float *, float *, float *, float *, cudaStream_t)
{ set_error("bf16 path not built yet"); return PCP_ERR_UNSUPPORTED; }
int model_prepare_bf16(pcp_model *) { set_error("bf16 path not built yet"); return PCP_ERR_UNSUPPORTED; }
void model_destroy_bf16(pcp_model *) {}
}
