// approx_gemm_c8.cu -- byte-cell instantiations of the table GEMM (lut_gemm.cuh, C8 = true), in a translation unit
// of their own so that they compile beside approx_gemm.cu.  See LutGemmArgs::c8 for what the variant does; it
// computes the bits of the 16-bit-cell kernel (tests/test_gpu_c8.py) and is chosen on the device per call.
#include "lut_gemm.cuh"

namespace b200 {

namespace {
struct C8Tile {
    int warps, rpw, cpl, kt, sidx;
    void (*kernel)(LutGemmArgs);
    void (*sk_kernel)(LutGemmArgs);
};
#define C8_TILE(W, R, C, KT, S) \
    {W, R, C, KT, S, approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, false, false, true>, \
     approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, false, true, true>}
// the tile list of approx_gemm.cu (kLutTiles)
const C8Tile kC8Tiles[] = {
    C8_TILE(8, 16, 4, 16, 0), C8_TILE(8, 32, 2, 16, 0), C8_TILE(4, 16, 4, 16, 0), C8_TILE(4, 32, 2, 16, 0),
    C8_TILE(4, 16, 2, 16, 0), C8_TILE(16, 16, 4, 16, 0),
    C8_TILE(16, 16, 4, 8, 1), C8_TILE(8, 16, 4, 8, 1), C8_TILE(8, 16, 4, 16, 1), C8_TILE(4, 16, 4, 16, 1),
    C8_TILE(4, 16, 2, 16, 1),
};
}  // namespace

bool lut_c8_kernels(int warps, int rpw, int cpl, int kt, int sidx, void (**kernel)(LutGemmArgs), void (**sk_kernel)(LutGemmArgs))
{
    for (const C8Tile& t : kC8Tiles)
        if (t.warps == warps && t.rpw == rpw && t.cpl == cpl && t.kt == kt && t.sidx == sidx) {
            *kernel = t.kernel;
            *sk_kernel = t.sk_kernel;
            return true;
        }
    return false;
}

int lut_c8_set_attributes()
{
    for (const C8Tile& t : kC8Tiles) {
        B200_CUDA_TRY(cudaFuncSetAttribute(t.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(t.sk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    }
    return B200_OK;
}

}  // namespace b200
