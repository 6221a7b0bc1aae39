// gemm_tc5_nt128.cu -- one configuration of the CTA-pair kernel: token tile, 128-k sub-blocks per shared-memory stage and per A
// stage, A stages in TMEM
#include "gemm_tc5.cuh"
namespace vlut {
namespace tc5 {
VLUT_TC5_INSTANCE(launch_nt128_a2, 128, 4, 4, 2)
}  // namespace tc5
}  // namespace vlut
