// weak fallbacks: overridden by ps_worker_tc.cu / ps_eval_tc.cu when those are linked in
#include "comm.cuh"
__attribute__((weak)) int psb_launch_worker_tc(const PsbSolver*, const CommCtx&, int, const void*, const int*, const int*, const float*,
                         float*, float*, int*, float*, cudaStream_t) {
    psb_set_error("tcgen05 worker kernel is not built into this library");
    return PSB_E_INVALID;
}
__attribute__((weak)) int psb_launch_eval_tc(const void*, const int*, int, const float*, int, int, int*, cudaStream_t) {
    psb_set_error("tcgen05 evaluation kernel is not built into this library");
    return PSB_E_INVALID;
}
