// epilogue.cuh -- SURVEY.md 8(f) N2: what the reference does to a layer's raw product sums before the next layer
// sees them, evaluated where the sums are produced instead of in separate passes over the output:
//   v = acc * u                       the GEMM's own scaling (exact power of two)
//   v = v + bias[c]                   `output += bias` (approx/approx_calculation.py:1020-1021 Linear,
//                                      :911-912 conv), torch's fp32 add
//   v = fma(v, alpha[c], beta'[c])    F.batch_norm in eval mode (quantization/quantized_folded_bn.py:66-72) in the op
//                                      order of ATen's CPU kernel: invstd = 1/sqrt(var + eps), alpha = gamma * invstd,
//                                      beta' = fma(-mean, alpha, beta) -- what b200_bn_act evaluates as its own pass
//   v = max(v, 0) [min(v, 6)]         the layer's ReLU / ReLU6 (quantized_folded_bn.py:74-75, hijacker.py:108-109)
//   store                             row-major [rows, N] or, for the conv form, contiguous NCHW -- the
//                                      `.view(B, Ho, Wo, C).permute(0, 3, 1, 2)` of approx_calculation.py:909-910 made
//                                      real (the separate pass did this transposition as well)
// Same operations in the same order as the unfused sequence: bit-identical results.  Off (on == 0) unless the caller
// passes a b200_epilogue: the default pass structure is the reference's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

struct Epilogue {
    int on;
    const float* bias;                          // [Ctot] or null
    const float *mean, *var, *gamma, *beta;     // eval BN statistics / affine ([Ctot]); mean == null: no BN
    float eps;
    int act;                                    // 0 none, 1 ReLU, 2 ReLU6
    int64_t HW;                                 // > 0: NCHW store, row m = b * HW + pixel;  0: row-major C[m * ldc + n]
    int Ctot, c0;                               // channels of the output tensor / first channel of this GEMM (group)

    struct Col {
        float bias, alpha, beta;
    };
    // per-output-channel constants (n = column of this GEMM)
    __device__ __forceinline__ Col column(int64_t n) const
    {
        Col c{0.0f, 1.0f, 0.0f};
        const int64_t ch = c0 + n;
        if (bias) c.bias = __ldg(bias + ch);
        if (mean) {
            const float invstd = 1.0f / sqrtf(__ldg(var + ch) + eps);
            c.alpha = (gamma ? __ldg(gamma + ch) : 1.0f) * invstd;
            c.beta = fmaf(-__ldg(mean + ch), c.alpha, beta ? __ldg(beta + ch) : 0.0f);
        }
        return c;
    }
    __device__ __forceinline__ float apply(float v, const Col& c) const
    {
        if (bias) v = v + c.bias;
        if (mean) v = fmaf(v, c.alpha, c.beta);
        if (act >= 1) v = fmaxf(v, 0.0f);
        if (act == 2) v = fminf(v, 6.0f);
        return v;
    }
    // element offset of (row m, column n) in the output tensor
    __device__ __forceinline__ int64_t index(int64_t m, int64_t n, int64_t ldc) const
    {
        if (HW <= 0) return m * ldc + n;
        const int64_t b = m / HW, p = m - b * HW;
        return (b * Ctot + c0 + n) * HW + p;
    }
    // generic (slow-path) store used by the fallback kernels: one element, constants computed on the spot
    __device__ __forceinline__ void store(float* C, int64_t ldc, int64_t m, int64_t n, float v) const
    {
        if (!on) { C[m * ldc + n] = v; return; }
        C[index(m, n, ldc)] = apply(v, column(n));
    }
};

inline Epilogue no_epilogue()
{
    Epilogue e{};
    return e;
}

}  // namespace b200
