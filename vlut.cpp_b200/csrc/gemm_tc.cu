// gemm_tc.cu -- VLUT_ALGO_TC placeholder (filled in below in the build order)
#include "vlut_common.cuh"
namespace vlut {
bool gemm_tc_supported(int, int64_t, int64_t, int64_t) { return false; }
cudaError_t launch_gemm_tc(const GemmArgs &, cudaStream_t, int *) { return cudaErrorNotSupported; }
}  // namespace vlut
