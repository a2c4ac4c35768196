// tcgen05 (UMMA) path for the res* models.  Placeholder until the kernel lands: no plan is
// built, so honk_model_create keeps the FP32 CUDA-core path.
#include "model.cuh"

namespace honk {
struct ResTcPlan {};
void res_tc_plan_free(ResTcPlan* p) { delete p; }
int res_tc_plan_create(honk_model* m) { m->tc = nullptr; return HONK_OK; }
int64_t res_tc_workspace_bytes(const honk_model*, int64_t) { return 0; }
int res_tc_forward(const honk_model*, const float*, int64_t, float*, void*, int64_t, cudaStream_t) {
  return set_error(HONK_ERR_UNSUPPORTED, "tcgen05 path not built");
}
}  // namespace honk
