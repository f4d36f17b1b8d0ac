// Evaluation, tensor-core path (tcgen05 + TMA + TMEM) -- placeholder until the kernel lands.
#include "common.cuh"

namespace cmf {
int eval_tile_tc(const cmf_eval_t* a, cudaStream_t s) {
  (void)a; (void)s;
  return fail(CMF_ERR_UNSUPPORTED, "cmf_eval_tile: precision=1 (tcgen05) path not built yet");
}
}  // namespace cmf
