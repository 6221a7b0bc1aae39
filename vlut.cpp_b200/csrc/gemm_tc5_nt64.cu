// gemm_tc5_nt64.cu -- one configuration of the CTA-pair kernel: token tile, 128-k sub-blocks per shared-memory stage and per A
// stage, A stages in TMEM
#include "gemm_tc5.cuh"
namespace vlut {
namespace tc5 {
VLUT_TC5_INSTANCE(launch_nt64, 64, 2, 4, 2)
}  // namespace tc5
}  // namespace vlut
