// gemm_impl.cuh -- internal interface between the C-ABI entry points (approx_gemm.cu, conv.cu).
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "lut.cuh"
#include "pair_gemm.cuh"

namespace b200 {

// Exponent biases either come from the host (python ints) or stay on the device as the fp32 tensors
// the fake-quantizer produced (FPQuantizer.custom_bias): no .item() sync on the hot path.
struct BiasSrc {
    int bA, bR;             // used when fA / fR are null
    const float* fA;        // device [1] or null
    const float* fR;        // device [1] or null
    const int32_t* iB;      // device [N] int32, or null
    const float* fB;        // device fp32, element n at fB[n * sB] (cast like .to(torch.int32))
    int64_t sB;
    __device__ __forceinline__ int A() const { return fA ? (int)__ldg(fA) : bA; }
    __device__ __forceinline__ int R() const { return fR ? (int)__ldg(fR) : bR; }
    __device__ __forceinline__ int B(int64_t n) const { return iB ? __ldg(iB + n) : (int)__ldg(fB + n * sB); }
};


// fused input fake-quantizer of an operand view (approx_gemm.cu)
int apply_act_quant(ASrc* a, const b200_act_quant* aq);

// (device, host table) -> cached device copy of a 4^M int8 table; nullptr stays nullptr (zero table)
int device_table(const signed char* host, int M, const signed char** dev_out);

// pair table with the single-column ("tensor bias") numerics, signed-index layout (approx_gemm.cu)
int get_quirk_lut(int E, int M, const signed char* host_table, const signed char* dev_table, int with_approx,
                  LutGeom* geom, const uint16_t** d_lut, int* lut_bytes, int* usable);

int approx_gemm_impl(const ASrc& A, const float* B, int64_t sb_k, int64_t sb_n, float* C, int64_t ldc,
                     int64_t Mrows, int64_t K, int64_t N, const BiasSrc& bias, const b200_approx_params* p,
                     void* workspace, size_t workspace_bytes, void* stream, const Epilogue* ep = nullptr);

// b200_epilogue (C ABI) -> Epilogue; HW / Ctot / c0 are set by the caller (conv form).  B200_OK or an argument error.
int make_epilogue(const b200_epilogue* e, Epilogue* out);

// extra workspace (beyond the GEMM-form bytes) with which the conv form decodes the input once and lets the
// table GEMM do the im2col gather (approx_gemm.cu)
size_t conv_gather_workspace_bytes(int64_t nB, int64_t Cg, int64_t H, int64_t W, int64_t Kg);

// quantize_after_mult_and_add GEMM: table kernel when `hint` (operand exponent biases) is given and the
// format allows it, arithmetic kernel otherwise (approx_gemm.cu)
int qamaa_gemm_impl(const ASrc& A, const float* B, int64_t sb_k, int64_t sb_n, float* C, int64_t ldc, int64_t Mrows,
                    int64_t K, int64_t N, const float* maxval_dev, int n_bits, float mantissa_bits, int sign_bits,
                    const BiasSrc* hint, void* workspace, size_t workspace_bytes, void* stream);

}  // namespace b200
