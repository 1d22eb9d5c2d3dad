// qamaa_gemm.cu -- K2: the quantize_after_mult_and_add GEMM (second mode).
//
// Replaces the per-column Python loop approx/approx_calculation.py:959-967 (Linear) / :787-795
// (BN-conv):
//     C[m,n] = Q( sum_k Q( A[m,k] * B[k,n] ) ),   Q = quantize_to_fp8_ste_MM(., n_bits, res maxval,
//                                                   mantissa_bits, sign_bits)[0]
// (quantization/quantizers/fp8_quantizer.py:97-173), `maxval` being the RESULT quantizer's range.
// One launch for all columns; `maxval` is read on the device (no .item()).
//
// Q here is evaluated without log2f: floor(log2|x| + bias) is the fp32 exponent field plus bias.
// The reference's fp32 `log2(|x|) + bias` can round UP a binade when |x| is a few ulps below a
// power of two; both binades then round x to that power of two, so the result is identical
// (tests/test_gpu_parity.py::test_qamaa_* checks it against the oracle, which uses log2f).
#include "common.cuh"
#include "conv_geom.cuh"
#include "pair_gemm.cuh"
#include "pair_math.cuh"
#include "qamaa_op.cuh"
#include "gemm_impl.cuh"

using namespace b200;

extern "C" int b200_qamaa_gemm(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, float* C,
                               int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* maxval_dev, int n_bits,
                               float mantissa_bits, int sign_bits, void* stream)
{
    if (int rc = check_device()) return rc;
    if (Mrows < 0 || K < 0 || N < 0 || !maxval_dev) { set_error("b200_qamaa_gemm: bad argument"); return B200_ERR_ARGUMENT; }
    if (Mrows == 0 || N == 0) return B200_OK;
    if (!A || !B || !C) { set_error("b200_qamaa_gemm: NULL operand"); return B200_ERR_ARGUMENT; }
    return qamaa_gemm_impl(make_gemm_asrc(A, lda), B, sb_k, sb_n, C, ldc, Mrows, K, N, maxval_dev, n_bits, mantissa_bits,
                           sign_bits, nullptr, nullptr, 0, stream);
}

extern "C" int b200_qamaa_gemm_fbias(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, float* C,
                                     int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* maxval_dev,
                                     int n_bits, float mantissa_bits, int sign_bits, const float* bA_dev,
                                     const float* bB_dev, int64_t bB_stride, void* workspace, size_t workspace_bytes,
                                     void* stream)
{
    if (int rc = check_device()) return rc;
    if (Mrows < 0 || K < 0 || N < 0 || !maxval_dev) { set_error("b200_qamaa_gemm_fbias: bad argument"); return B200_ERR_ARGUMENT; }
    if (Mrows == 0 || N == 0) return B200_OK;
    if (!A || !B || !C) { set_error("b200_qamaa_gemm_fbias: NULL operand"); return B200_ERR_ARGUMENT; }
    const int E = n_bits - sign_bits - (int)rintf(mantissa_bits);
    const int dflt = 1 << (E > 0 ? E - 1 : 0);
    BiasSrc hint{dflt, dflt, bA_dev, nullptr, nullptr, bB_dev, bB_stride};
    return qamaa_gemm_impl(make_gemm_asrc(A, lda), B, sb_k, sb_n, C, ldc, Mrows, K, N, maxval_dev, n_bits, mantissa_bits,
                           sign_bits, bB_dev ? &hint : nullptr, workspace, workspace_bytes, stream);
}

// Conv form (BN-conv variant, approx/approx_calculation.py:787-795 inside the group loop :858-906):
// implicit im2col over the NCHW input, one launch per group.  out is [B*Ho*Wo, C_out] row-major.
// Groups with a single output channel take the reference's N == 1 branch, which has NO
// quantize_after_mult_and_add mode (:800-811): the caller must use the exact matmul for them.
static int qamaa_conv2d_impl(const float* x, const float* w, float* out, const b200_conv_geom* g, const float* maxval_dev,
                            int n_bits, float mantissa_bits, int sign_bits, const float* bA_dev, const float* bB_dev,
                            int64_t bB_stride, void* workspace, size_t workspace_bytes, void* stream)
{
    if (int rc = check_device()) return rc;
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    if (!maxval_dev) { set_error("b200_qamaa_conv2d: maxval is NULL"); return B200_ERR_ARGUMENT; }
    if (d.Ng == 1) {
        set_error("b200_qamaa_conv2d: single-output-channel groups have no quantize_after_mult_and_add mode in the reference");
        return B200_ERR_ARGUMENT;
    }
    if (d.rows == 0) return B200_OK;
    if (!x || !w || !out) { set_error("b200_qamaa_conv2d: NULL operand"); return B200_ERR_ARGUMENT; }
    const int E = n_bits - sign_bits - (int)rintf(mantissa_bits);
    const int dflt = 1 << (E > 0 ? E - 1 : 0);
    for (int grp = 0; grp < g->groups; ++grp) {
        const float* wg = w + (int64_t)grp * d.Ng * d.Kg;     // weight_col[g*Ng:(g+1)*Ng, :] ([N,K] rows, k contiguous)
        BiasSrc hint{dflt, dflt, bA_dev, nullptr, nullptr, bB_dev ? bB_dev + (int64_t)grp * d.Ng * bB_stride : nullptr,
                     bB_stride};
        if (int rc = qamaa_gemm_impl(make_conv_asrc(x, g, d, grp), wg, 1, d.Kg, out + (int64_t)grp * d.Ng, g->C_out,
                                     d.rows, d.Kg, d.Ng, maxval_dev, n_bits, mantissa_bits, sign_bits,
                                     bB_dev ? &hint : nullptr, workspace, workspace_bytes, stream))
            return rc;
    }
    return B200_OK;
}

extern "C" int b200_qamaa_conv2d(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                 const float* maxval_dev, int n_bits, float mantissa_bits, int sign_bits, void* stream)
{
    return qamaa_conv2d_impl(x, w, out, g, maxval_dev, n_bits, mantissa_bits, sign_bits, nullptr, nullptr, 0, nullptr, 0,
                             stream);
}

extern "C" size_t b200_qamaa_conv2d_workspace_bytes(const b200_conv_geom* g)
{
    ConvDims d;
    if (conv_dims(g, &d) != B200_OK) return 0;
    return b200_qamaa_gemm_workspace_bytes(d.rows, d.Kg, d.Ng) + conv_gather_workspace_bytes(g->batch, d.Cg, g->H, g->W, d.Kg);
}

extern "C" int b200_qamaa_conv2d_fbias(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                       const float* maxval_dev, int n_bits, float mantissa_bits, int sign_bits,
                                       const float* bA_dev, const float* bB_dev, int64_t bB_stride, void* workspace,
                                       size_t workspace_bytes, void* stream)
{
    return qamaa_conv2d_impl(x, w, out, g, maxval_dev, n_bits, mantissa_bits, sign_bits, bA_dev, bB_dev, bB_stride,
                             workspace, workspace_bytes, stream);
}
