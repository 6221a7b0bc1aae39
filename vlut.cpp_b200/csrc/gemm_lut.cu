// gemm_lut.cu -- VLUT_ALGO_LUT placeholder (filled in below in the build order)
#include "vlut_common.cuh"
namespace vlut {
bool gemm_lut_supported(int, int64_t, int64_t, int64_t) { return false; }
cudaError_t launch_gemm_lut(const GemmArgs &, cudaStream_t, int *) { return cudaErrorNotSupported; }
}  // namespace vlut
