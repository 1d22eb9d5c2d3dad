// qamaa_op.cuh -- the quantize_after_mult_and_add pair operator (quantize_to_fp8_ste_MM applied to every
// product and to the sum; approx/approx_calculation.py:959-967, quantization/quantizers/fp8_quantizer.py:97-173),
// shared by the arithmetic kernel (qamaa_gemm.cu) and the table path (approx_gemm.cu).
#pragma once
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
        const float xc = clamp_nan(x, mn, mx);                             // :119-120
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
