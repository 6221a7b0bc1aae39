// gemm_tc5.cu -- host-side dispatch of VLUT_ALGO_TCP (the CTA-pair kernel, gemm_tc5.cuh) over its token-tile configurations.
#include "gemm_tc5.cuh"

namespace vlut {

namespace tc5 {
cudaError_t launch_nt64(Params &p, int sm_count, cudaStream_t s);        // gemm_tc5_nt64.cu
cudaError_t launch_nt128_a2(Params &p, int sm_count, cudaStream_t s);    // gemm_tc5_nt128.cu
cudaError_t launch_nt128_a4(Params &p, int sm_count, cudaStream_t s);    // gemm_tc5_nt128b.cu
cudaError_t launch_nt128_a3(Params &p, int sm_count, cudaStream_t s);    // gemm_tc5_nt128c.cu
cudaError_t launch_nt256_a2(Params &p, int sm_count, cudaStream_t s);    // gemm_tc5_nt256.cu
cudaError_t launch_nt256_a4(Params &p, int sm_count, cudaStream_t s);    // gemm_tc5_nt256b.cu
}  // namespace tc5

bool gemm_tc5_supported(int type, int64_t K, int64_t N, int64_t T) {
    return type_valid(type) && T > 32 && K >= 4 && N >= 1 && ceil_div(K, group_size(type)) <= 64;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (the library links cudart only)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn) f;
    }
    return fn;
}
// out [T][ldo] f32 as a 2-D tensor {N, T}, box {128 rows, 16 tokens}: the staged drain's TMA stores (OUT == 4)
static bool make_out_map(CUtensorMap *m, float *out, int64_t N, int64_t T, int64_t ldo) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn) return false;
    const cuuint64_t dims[2] = {(cuuint64_t) N, (cuuint64_t) T};
    const cuuint64_t strides[1] = {(cuuint64_t) ldo * 4};
    const cuuint32_t box[2] = {128, 16};
    const cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

cudaError_t launch_gemm_tc5(const GemmArgs &g, cudaStream_t s, int *launches) {
    if (g.act.sw_tile * 2 != g.act.NT) return cudaErrorInvalidValue;  // wants half-height swizzled activation tiles
    tc5::Params p{};
    p.g = g;
    p.KB = (int) (g.act.Kpad / 128);
    p.gsub = g.act.gsize / 128;
    p.G = g.act.G;
    p.RT = (int) (g.Npad / TILE_ROWS);
    p.RP = (p.RT + 1) / 2;
    const int NT = g.act.NT;
    p.TT = (int) (g.act.Tpad / NT);
    p.U = p.RP * p.TT * p.G;
    const Knobs &kn = knobs();
    if (p.g.local_bulk && !make_out_map(&p.out_map, g.out, g.N, g.T, g.ldo)) p.g.local_bulk = 0;   // no tensor map: direct stores
    // local output: 16-byte stores (four rows of a token per lane after a 4 x 4 transpose inside the warp) when the rows allow it
    if (g.n_out == 0) p.g.vec_ok = (kn.tc_vec_out && g.ldo % 4 == 0 && (reinterpret_cast<uintptr_t>(g.out) & 15) == 0) ? 1 : 0;
    p.allow_fast = kn.tc_fast;
    p.uniform = (kn.tc_lean && p.KB % p.gsub == 0) ? 1 : 0;
    p.pdl = kn.pdl;
    p.abl = kn.tc_abl;
    p.dbg_skip = kn.tc_debug_skip;
    p.dbg_cta = kn.tcs_dbg_cta > 0 ? kn.tcs_dbg_cta : 0;   // VLUT_TCS_DBG_CTA picks the recorded CTA here too
    p.dbg = kn.tc_debug ? debug_buffer(s, kn.tc_debug_skip) : nullptr;
    if (launches) *launches += 1;
    switch (NT) {
        case 64: return tc5::launch_nt64(p, g.sm_count, s);
        // A-stage depth (measured, profiles/r2_tcp_tiles.txt): the 128-token tile is faster with two hand-offs of 512 k
        // (T = 512: 26.8 vs 32.5 us), the 256-token tile with four of 256 k (T = 2048: 87.0 vs 91.1 us)
        case 128: return kn.tc_astages == 4 ? tc5::launch_nt128_a4(p, g.sm_count, s)
                       : kn.tc_astages == 3 ? tc5::launch_nt128_a3(p, g.sm_count, s) : tc5::launch_nt128_a2(p, g.sm_count, s);
        case 256: return kn.tc_astages == 2 ? tc5::launch_nt256_a2(p, g.sm_count, s) : tc5::launch_nt256_a4(p, g.sm_count, s);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace vlut
