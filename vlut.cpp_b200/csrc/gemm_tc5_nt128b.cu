// gemm_tc5_nt128b.cu -- one configuration of the CTA-pair kernel: token tile, 128-k sub-blocks per shared-memory stage and per A
// stage, A stages in TMEM
#include "gemm_tc5.cuh"
namespace vlut {
namespace tc5 {
VLUT_TC5_INSTANCE(launch_nt128_a4, 128, 2, 2, 4)
}  // namespace tc5
}  // namespace vlut
