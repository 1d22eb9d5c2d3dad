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

namespace b200 {

struct QamaaOp {
    struct Col {};
    const float* maxval_dev;
    int n_bits, sign_bits;
    float mantissa_bits;
    // derived per thread
    float mx, mn, bias, Mf;
    int bias_i, M_i, fast;
    __device__ __forceinline__ void begin()
    {
        mx = __ldg(maxval_dev);
        mn = (sign_bits == 1) ? -mx : 0.0f;
        bias = fp8_ste_bias(mx, n_bits, mantissa_bits, sign_bits, &Mf);
        bias_i = (int)bias;
        M_i = (int)Mf;
        // exponent-field path needs an integral, moderate bias (always true for a finite positive maxval)
        fast = (bias == (float)bias_i) && bias_i > -90 && bias_i < 90;
    }
    __device__ __forceinline__ Col column(int64_t) const { return Col{}; }
    __device__ __forceinline__ float quant(float x) const
    {
        if (!fast) return fp8_ste_quant(x, mx, bias, Mf, sign_bits);
        const float xc = fminf(fmaxf(x, mn), mx);                          // :119-120
        const int ef = (int)((__float_as_uint(xc) >> 23) & 0xFFu);          // floor(log2|xc|) + 127 (0: zero / denormal)
        int ls = ef - 127 + bias_i;                                         // :144
        if (ls < 1) ls = 1;                                                 // clamp(min=1)
        const int se = ls - M_i - bias_i;                                   // log2(scale), :151
        // round-half-even to a multiple of 2^se: add and subtract 1.5 * 2^(23+se) (exact; |xc| < 2^(22+se))
        const float magic = __uint_as_float(((uint32_t)(se + 23 + 127) << 23) | 0x00400000u);
        return (xc + magic) - magic;                                        // == rint(xc / scale) * scale, :154
    }
    __device__ __forceinline__ float pair(float a, float b, const Col&) const { return quant(a * b); }
    __device__ __forceinline__ float finish(float acc, const Col&) const { return quant(acc); }
};

}  // namespace b200

using namespace b200;

extern "C" int b200_qamaa_gemm(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, float* C,
                               int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* maxval_dev, int n_bits,
                               float mantissa_bits, int sign_bits, void* stream)
{
    if (int rc = check_device()) return rc;
    if (Mrows < 0 || K < 0 || N < 0 || !maxval_dev) { set_error("b200_qamaa_gemm: bad argument"); return B200_ERR_ARGUMENT; }
    if (Mrows == 0 || N == 0) return B200_OK;
    if (!A || !B || !C) { set_error("b200_qamaa_gemm: NULL operand"); return B200_ERR_ARGUMENT; }
    QamaaOp op;
    op.maxval_dev = maxval_dev;
    op.n_bits = n_bits;
    op.sign_bits = sign_bits;
    op.mantissa_bits = mantissa_bits;
    pair_gemm_kernel<QamaaOp><<<pair_gemm_grid(Mrows, N), 256, 0, (cudaStream_t)stream>>>(
        make_gemm_asrc(A, lda), BSrc{B, sb_k, sb_n}, C, ldc, Mrows, K, N, op, nullptr, 0);
    B200_LAUNCH_CHECK("pair_gemm_kernel<QamaaOp>");
    return B200_OK;
}

// Conv form (BN-conv variant, approx/approx_calculation.py:787-795 inside the group loop :858-906):
// implicit im2col over the NCHW input, one launch per group.  out is [B*Ho*Wo, C_out] row-major.
// Groups with a single output channel take the reference's N == 1 branch, which has NO
// quantize_after_mult_and_add mode (:800-811): the caller must use the exact matmul for them.
extern "C" int b200_qamaa_conv2d(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                 const float* maxval_dev, int n_bits, float mantissa_bits, int sign_bits, void* stream)
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
    QamaaOp op;
    op.maxval_dev = maxval_dev;
    op.n_bits = n_bits;
    op.sign_bits = sign_bits;
    op.mantissa_bits = mantissa_bits;
    for (int grp = 0; grp < g->groups; ++grp) {
        const float* wg = w + (int64_t)grp * d.Ng * d.Kg;     // weight_col[g*Ng:(g+1)*Ng, :] ([N,K] rows, k contiguous)
        pair_gemm_kernel<QamaaOp><<<pair_gemm_grid(d.rows, d.Ng), 256, 0, (cudaStream_t)stream>>>(
            make_conv_asrc(x, g, d, grp), BSrc{wg, 1, d.Kg}, out + (int64_t)grp * d.Ng, g->C_out, d.rows, d.Kg, d.Ng,
            op, nullptr, 0);
        B200_LAUNCH_CHECK("pair_gemm_kernel<QamaaOp>(conv)");
    }
    return B200_OK;
}
