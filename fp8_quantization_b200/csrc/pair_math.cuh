// pair_math.cuh -- per-element / per-pair arithmetic of the approx_v9 path, device side.
//
// This is the arithmetic definition the generic kernel executes per operand pair and the table
// builder evaluates once per (t, mA, mB) cell.  It follows the reference's torch op chain in fp32
// (file:line cited per function; paths relative to the reference root).  The library is compiled
// with -fmad=false so no multiply-add is fused behind our back.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace b200 {

// 2.0 ** int32 tensor in fp32 (exact power of two, inf above the range, 0 far below).
__host__ __device__ __forceinline__ float pow2f_i(int n)
{
    if (n > 127) return INFINITY;
    if (n < -149) return 0.0f;
    return ldexpf(1.0f, n);
}

// 2 ** int32 tensor (integer pow): a negative exponent gives 0 -- this is what switches
// sub-normal handling off in the single-column branch.
__host__ __device__ __forceinline__ int ipow2_i(int n)
{
    return (n < 0 || n >= 32) ? 0 : (int)(1u << n);
}

// param_prepare, approx/approx_matmul_whole_v9.py:189-229
struct FmtParam {
    int M, bias, max_expo, max_mant;
    float mant_scale, max_norm, min_norm;
};

__host__ __device__ inline FmtParam make_param(int E, int M, int bias, bool tensor_bias_quirk)
{
    FmtParam p;
    p.M = M;
    p.bias = bias;
    p.max_expo = (1 << E) - 1;
    p.max_mant = (1 << M) - 1;
    p.mant_scale = (float)(1 << M);
    const float top = 2.0f - ldexpf(1.0f, -M);
    if (!tensor_bias_quirk) {
        p.max_norm = pow2f_i(p.max_expo - bias) * top;   // :201 (python numbers, exact)
        p.min_norm = pow2f_i(1 - bias);                  // :202
    } else {
        p.max_norm = (float)ipow2_i(p.max_expo - bias) * top;  // integer pow on int32 tensors
        p.min_norm = (float)ipow2_i(1 - bias);
    }
    return p;
}

// float_to_fpany_absint_torch, approx/approx_matmul_whole_v9.py:233-291
__host__ __device__ inline void fpany_encode(const FmtParam& p, float v, bool clip_OF, int& expo, int& mant)
{
    int ex = 0;
    float m = frexpf(v, &ex);                                   // :261
    if (v == 0.0f) { m = 0.0f; ex = 0; }
    const bool sub = fabsf(v) < p.min_norm;                     // :264
    float x;
    if (sub) x = fabsf(m) * pow2f_i(ex + p.bias - 1 + p.M);     // :269,:274
    else     x = (fabsf(m) * 2.0f - 1.0f) * pow2f_i(p.M);       // :275
    float r = rintf(x);                                         // torch.round = half to even
    if (r > (float)p.max_mant) r = (float)p.max_mant;           // clamp(max=max_mant), no carry
    mant = (int)r;
    expo = sub ? 0 : ex + (p.bias - 1);                         // :280
    if (clip_OF && (v < -p.max_norm || v > p.max_norm)) {       // :283-286
        expo = p.max_expo;
        mant = p.max_mant;
    }
}

// fpany_absint_to_float_torch, approx/approx_matmul_whole_v9.py:295-329
__host__ __device__ inline float fpany_decode(const FmtParam& p, float sign, int expo, int mant)
{
    const float frac = (float)mant / p.mant_scale;
    const float v = (expo == 0) ? pow2f_i(1 - p.bias) * frac               // :321-323
                                : pow2f_i(expo - p.bias) * (1.0f + frac);
    return v * sign;
}

// quant_to_fp_any_vectorize_torch, approx/approx_matmul_whole_v9.py:333-362
__host__ __device__ inline float fpany_quant(const FmtParam& p, float v, bool clip_OF)
{
    int e, m;
    fpany_encode(p, v, clip_OF, e, m);
    return fpany_decode(p, (v < 0.0f) ? -1.0f : 1.0f, e, m);               // :357
}

// all-normal codec, approx/fp_any_int_v5.py:6-53, :55-110, :116-162
__host__ __device__ inline void allnorm_encode(int E, int M, int bias, float v, bool clip_OF, int& expo, int& mant)
{
    const int max_expo = (1 << E) - 1, max_mant = (1 << M) - 1;
    const float max_value = pow2f_i(max_expo - bias) * (2.0f - ldexpf(1.0f, -M));
    const float min_value = pow2f_i(-bias) * (1.0f + ldexpf(1.0f, -M));
    int ex = 0;
    float m = frexpf(v, &ex);
    if (v == 0.0f) { m = 0.0f; ex = 0; }
    const bool zero = (v > -min_value) && (v < min_value);                 // :92
    float r = rintf(zero ? 0.0f : (fabsf(m) * 2.0f - 1.0f) * pow2f_i(M));  // :95-99
    if (r > (float)max_mant) r = (float)max_mant;
    mant = (int)r;
    expo = zero ? 0 : ex + (bias - 1);                                     // :102
    if (clip_OF && (v < -max_value || v > max_value)) { expo = max_expo; mant = max_mant; }
}

__host__ __device__ inline float allnorm_decode(int M, int bias, float sign, int expo, int mant)
{
    if (expo == 0 && mant == 0) return 0.0f * sign;                        // :153-156
    return pow2f_i(expo - bias) * (1.0f + (float)mant / (float)(1 << M)) * sign;
}

struct PairCfg {
    int E, M;
    bool with_approx, s2nn2s, clip_OF, qbma, quirk;
    // ofuf_mode 1: the v5 integer-add GEMM (pair_v5 below) with its live overflow / underflow flags
    bool v5 = false, sim_ofuf = false, of_opt = false, uf_opt = false;
};

// mult_result_mant, approx/approx_matmul_whole_v9.py:178-184 (negative indices wrap like python)
template <typename TabT>
__host__ __device__ inline float mult_mant(int M, int mA, int mB, const TabT* T, bool with_approx)
{
    const float s = ldexpf(1.0f, -M);
    const float prod = (1.0f + (float)mA * s) * (1.0f + (float)mB * s);
    if (!with_approx) return prod;
    const int n = 1 << M;
    const int ia = mA < 0 ? mA + n : mA, ib = mB < 0 ? mB + n : mB;
    const float t = (ia >= 0 && ia < n && ib >= 0 && ib < n) ? (float)T[ia * n + ib] : NAN;
    return prod - s * t;
}

// One operand pair of custom_matmul_vectorize, approx/approx_matmul_whole_v9.py:29-110.
template <typename TabT>
__host__ __device__ inline float pair_generic(const PairCfg& c, const FmtParam& pA, const FmtParam& pB,
                                              const FmtParam& pR, const TabT* T, float a, float b)
{
    float g = a * b;                                            // :30-31
    const bool zero = (g == 0.0f);                              // :32
    if (c.qbma) g = fpany_quant(pR, g, c.clip_OF);              // :35-36
    const bool subA = fabsf(a) < pA.min_norm;                   // :47
    const bool subB = fabsf(b) < pB.min_norm;                   // :48
    float a2 = a, b2 = b;
    if (c.s2nn2s) {                                             // :52-54
        if (subA) a2 = a * pA.mant_scale;
        if (subB) b2 = b * pB.mant_scale;
    }
    int eA, mA, eB, mB;
    fpany_encode(pA, a2, false, eA, mA);                        // :58
    fpany_encode(pB, b2, false, eB, mB);                        // :59
    const int e = eA + eB - (pA.bias + pB.bias - pR.bias);      // :66-67
    const float sgn = (g < 0.0f) ? -1.0f : 1.0f;                // :68
    float r = pow2f_i(e - pR.bias) * mult_mant(c.M, mA, mB, T, c.with_approx) * sgn;  // :73-77 / :92-96
    if (c.s2nn2s) {
        if (subA) r = r / pA.mant_scale;                        // :79
        if (subB) r = r / pB.mant_scale;                        // :80
        if (zero) r = 0.0f;                                     // :81
    } else {
        const bool norm = (eA > 0) && (eB > 0) && (fabsf(g) >= pR.min_norm);  // :86-87
        if (!norm) r = g;                                       // :89-98
    }
    if (c.qbma) r = fpany_quant(pR, r, c.clip_OF);              // :107-108
    return r;
}

// torch.min(torch.max(x, minval), maxval) (fp8_quantizer.py:120): torch's min / max PROPAGATE NaN, fminf / fmaxf do not
__host__ __device__ __forceinline__ float clamp_nan(float x, float mn, float mx)
{
    return (x != x) ? x : fminf(fmaxf(x, mn), mx);
}

// One operand pair of the v5 GEMM (opt-in): custom_matmul_vectorize of
// approx/approx_matmul_whole_v5.py:10-152 with approx_mult_new :155-182.  T: v5 comp table or nullptr (zeros).
// One custom_bias in v5; per-operand biases enter as B_neg = -((bA + bB - bR) << M) (approx_v9's exponent formula).
template <typename TabT>
__host__ __device__ inline float pair_v5(const PairCfg& c, const FmtParam& pA, const FmtParam& pB, const FmtParam& pR,
                                         const TabT* T, float a, float b)
{
    int eA, mA, eB, mB;
    fpany_encode(pA, a, true, eA, mA);                          // :26-37, clip_OF = True (:22)
    fpany_encode(pB, b, true, eB, mB);                          // :39-50
    const int ms = 1 << c.M, mod = 1 << (c.E + c.M);
    int t = (eA * ms + mA) + (eB * ms + mB) - (pA.bias + pB.bias - pR.bias) * ms;   // :52-59, :159
    if (T) t += (int)T[mA * ms + mB];                           // :159 comp_table_NN[x_mant, y_mant]
    if (c.sim_ofuf) {                                           // :162-176
        const bool of = t > mod - 1, uf = t < 0;
        t = ((t % mod) + mod) % mod;                            // floor-mod
        if (c.of_opt && of) t = mod - 1;
        if (c.uf_opt && uf) t = t % ms;
    }
    const int expo = t >> c.M, mant = t & (ms - 1);             // :298-299 (floor division, non-negative remainder)
    const float sign = ((a < 0.0f) ? -1.0f : 1.0f) * ((b < 0.0f) ? -1.0f : 1.0f);   // :106-108
    return fpany_decode(pR, sign, expo, mant);                  // :111-114
}

template <typename TabT>
__host__ __device__ inline float pair_any(const PairCfg& c, const FmtParam& pA, const FmtParam& pB, const FmtParam& pR,
                                          const TabT* T, float a, float b)
{
    return c.v5 ? pair_v5(c, pA, pB, pR, T, a, b) : pair_generic(c, pA, pB, pR, T, a, b);
}

// quantize_to_fp8_ste_MM element body, quantization/quantizers/fp8_quantizer.py:119-154
__host__ __device__ inline float fp8_ste_quant(float x, float maxval, float bias, float Mf, int sign_bits)
{
    const float mn = (sign_bits == 1) ? -maxval : 0.0f;         // :119
    const float xc = clamp_nan(x, mn, maxval);                  // :120
    float ls = floorf(log2f(fabsf(xc)) + bias);                 // :144
    if (!(ls >= 1.0f)) ls = 1.0f;                               // clamp(min=1.0); -inf -> 1
    const float scale = pow2f_i((int)(ls - Mf - bias));         // :151 (integer-valued argument)
    return rintf(xc / scale) * scale;                           // :154
}

// The same element function without log2f / division, for an integral, moderate bias (what fp8_ste_bias returns
// for any finite positive maxval): floor(log2|xc|) is the exponent field, and round-half-even to a multiple of
// scale = 2^se is (xc + 1.5 * 2^(23+se)) - 1.5 * 2^(23+se), exact because |xc| < 2^(22+se).  Where log2f rounds
// up to an integer just below a power of two the reference picks the coarser scale and still lands on the same
// value (the all-ones mantissa rounds to the power of two on either grid).  Bit-identical to fp8_ste_quant on
// every reference fixture (tests/golden/fakequant.npz) and against the oracle's log2f form on raw operands.
__host__ __device__ inline bool fp8_ste_fast_ok(float bias) { const int b = (int)bias; return bias == (float)b && b > -90 && b < 90; }
__device__ __forceinline__ float fp8_ste_quant_fast(float x, float maxval, int bias_i, int M_i, int sign_bits)
{
    const float mn = (sign_bits == 1) ? -maxval : 0.0f;                 // :119
    const float xc = clamp_nan(x, mn, maxval);                          // :120
    const int ef = (int)((__float_as_uint(xc) >> 23) & 0xFFu);          // floor(log2|xc|) + 127 (0: zero / denormal)
    int ls = ef - 127 + bias_i;                                         // :144
    if (ls < 1) ls = 1;                                                 // clamp(min=1)
    const int se = ls - M_i - bias_i;                                   // log2(scale), :151
    const float magic = __uint_as_float(((uint32_t)(se + 23 + 127) << 23) | 0x00400000u);
    return (xc + magic) - magic;                                        // == rint(xc / scale) * scale, :154
}

// bias = round(2**E - log2(maxval) + log2(2 - 2**-M) - 1), fp8_quantizer.py:111-117, fp32 op by op
__host__ __device__ inline float fp8_ste_bias(float maxval, int n_bits, float mantissa_bits, int sign_bits, float* M_out)
{
    float Mf = rintf(mantissa_bits);
    Mf = fminf(fmaxf(Mf, 1.0f), (float)(n_bits - sign_bits));
    const float Ef = (float)(n_bits - sign_bits) - Mf;
    const float t1 = pow2f_i((int)Ef) - log2f(maxval);
    const float t2 = t1 + log2f(2.0f - pow2f_i(-(int)Mf));
    if (M_out) *M_out = Mf;
    return rintf(t2 - 1.0f);
}

}  // namespace b200
