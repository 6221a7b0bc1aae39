// gemm_tc5_nt256b.cu -- one configuration of the CTA-pair kernel: token tile, 128-k sub-blocks per shared-memory stage and per A
// stage, A stages in TMEM
#include "gemm_tc5.cuh"
namespace vlut {
namespace tc5 {
VLUT_TC5_INSTANCE(launch_nt256_a4, 256, 2, 2, 4)
}  // namespace tc5
}  // namespace vlut
