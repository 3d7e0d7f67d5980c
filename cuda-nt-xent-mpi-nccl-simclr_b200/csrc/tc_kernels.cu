// tc_kernels.cu -- placeholder until the tcgen05 kernels land (phase B).
#include "tc_kernels.cuh"
namespace ntxent_b200 {
bool tc_supported(int64_t, bool) { return false; }
SplitPlan tc_plan_fwd(int64_t, int64_t, int64_t, int) { return SplitPlan{}; }
SplitPlan tc_plan_bwd(int64_t, int64_t, int64_t, int64_t, int) { return SplitPlan{}; }
cudaError_t tc_launch_fwd(const void*, int64_t, int64_t, int64_t, int64_t, float, const SplitPlan&,
                          float*, float*, unsigned*, cudaStream_t) { return cudaErrorNotSupported; }
cudaError_t tc_launch_bwd(const void*, const float*, int64_t, int64_t, int64_t, int64_t, int64_t,
                          int64_t, float, int, const SplitPlan&, float*, cudaStream_t) {
    return cudaErrorNotSupported;
}
}  // namespace ntxent_b200
