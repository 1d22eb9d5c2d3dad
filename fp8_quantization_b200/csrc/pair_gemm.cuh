// pair_gemm.cuh -- register-blocked CUDA-core GEMM skeleton for "sum over k of f(a, b)" kernels whose
// pair function f is evaluated arithmetically (no table): the arithmetic approx_v9 kernel
// (approx_gemm.cu) and the quantize-after-mult-and-add kernel (qamaa_gemm.cu).
//
// CTA tile 64 x 64, K tile 16, 256 threads, 4 x 4 outputs per thread.  Each output is accumulated in
// fp32 by ONE thread with k ascending -- the oracle's order.  Operand tiles are staged k-major in
// shared memory so that a thread reads its 4 rows / 4 columns as one 128-bit LDS each.
//
// Operand sources are small functors so that the same kernel serves the GEMM form (A row-major,
// B any strided view such as weight.t()) and the conv form (implicit im2col over an NCHW tensor,
// reference: approx/approx_calculation.py:724-747).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "epilogue.cuh"
#include "pair_math.cuh"

namespace b200 {

// A operand: element (m, k) of the [Mrows, K] left matrix.
struct ASrc {
    const float* p;
    int64_t lda;     // GEMM form: p[m * lda + k]
    int conv;        // 1: implicit im2col, K order (c, ky, kx), rows (b, ho, wo)
    int C, H, W, kh, kw, sh, sw, ph, pw, dh, dw, Ho, Wo;
    int c0;          // first input channel of the group being computed
    // Fused input fake-quantizer: when fq_maxval is set, every element read through this
    // view is quantize_to_fp8_ste_MM(x, maxval, ...) (quantization/quantizers/fp8_quantizer.py:119-154) of the
    // raw tensor -- the same element function the stand-alone quantizer kernel evaluates, with the bias that
    // kernel derived from the same maxval, so the bits are those of quantizing first.  Q(0) = 0: padding commutes.
    const float* fq_maxval;   // device [1] or null
    const float* fq_bias;     // device [1]: FPQuantizer.custom_bias
    float fq_Mf;              // clamp(round(mantissa_bits), 1, n_bits - sign_bits)
    int fq_sign;
    __device__ __forceinline__ float fq(float v) const
    {
        if (!fq_maxval) return v;
        const float mx = __ldg(fq_maxval), bias = __ldg(fq_bias);
        return fp8_ste_fast_ok(bias) ? fp8_ste_quant_fast(v, mx, (int)bias, (int)fq_Mf, fq_sign)
                                     : fp8_ste_quant(v, mx, bias, fq_Mf, fq_sign);
    }
    __device__ __forceinline__ float load(int64_t m, int64_t k) const { return fq(load_raw(m, k)); }
    __device__ __forceinline__ float load_raw(int64_t m, int64_t k) const
    {
        if (!conv) return p[m * lda + k];
        const int kx = (int)(k % kw);
        const int t = (int)(k / kw);
        const int ky = t % kh;
        const int c = t / kh + c0;
        const int wo = (int)(m % Wo);
        const int64_t t2 = m / Wo;
        const int ho = (int)(t2 % Ho);
        const int64_t b = t2 / Ho;
        const int h = ho * sh - ph + ky * dh, w = wo * sw - pw + kx * dw;
        if (h < 0 || h >= H || w < 0 || w >= W) return 0.0f;    // F.pad zeros (:727)
        return p[((b * C + c) * H + h) * (int64_t)W + w];
    }
};

inline ASrc make_gemm_asrc(const float* A, int64_t lda)
{
    ASrc a{};
    a.p = A;
    a.lda = lda;
    a.conv = 0;
    return a;
}

// B operand: element (k, n) of the [K, N] right matrix, any strides.
struct BSrc {
    const float* p;
    int64_t sb_k, sb_n;
    __device__ __forceinline__ float load(int64_t k, int64_t n) const { return p[k * sb_k + n * sb_n]; }
};

constexpr int PG_BM = 64, PG_BN = 64, PG_KT = 16, PG_LD = 68;

// Op interface:
//   struct Col;                                   per-column state (e.g. the weight format of column n)
//   __device__ void  begin();                     once per thread
//   __device__ Col   column(int64_t n) const;     n < N guaranteed
//   __device__ float pair(float a, float b, const Col&) const;
//   __device__ float finish(float acc, const Col&) const;
template <class Op>
__global__ void __launch_bounds__(256) pair_gemm_kernel(ASrc A, BSrc B, float* __restrict__ C, int64_t ldc,
                                                        int64_t Mrows, int64_t K, int64_t N, Op op,
                                                        const int* __restrict__ gate, int gate_run_if_nonzero,
                                                        Epilogue ep = Epilogue{})
{
    if (gate) {
        const int g = *gate;
        if ((g != 0) != (gate_run_if_nonzero != 0)) return;
    }
    __shared__ __align__(16) float As[PG_KT][PG_LD];
    __shared__ __align__(16) float Bs[PG_KT][PG_LD];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    op.begin();
    // 1-D grid, tiles dealt round-robin (column tile fastest): the gated launches that accompany every table GEMM and
    // exit at once stay small however large the problem is
    const int64_t tiles_n = (N + PG_BN - 1) / PG_BN, tiles = ((Mrows + PG_BM - 1) / PG_BM) * tiles_n;
  for (int64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int64_t m0 = (tile / tiles_n) * PG_BM, n0 = (tile % tiles_n) * PG_BN;
    __syncthreads();   // the previous tile's readers of As / Bs are done
    typename Op::Col col[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t n = n0 + tx * 4 + j;
        col[j] = op.column(n < N ? n : (N - 1));
    }
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

    const bool b_k_contig = (B.sb_k == 1 && B.sb_n != 1);
    for (int64_t k0 = 0; k0 < K; k0 += PG_KT) {
        __syncthreads();
        {   // A tile: k fastest across threads (rows are k-contiguous in the GEMM form)
            const int kk = tid & 15;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int mm = (tid >> 4) + 16 * i;
                const int64_t m = m0 + mm, k = k0 + kk;
                As[kk][mm] = (m < Mrows && k < K) ? A.load(m, k) : 0.0f;
            }
        }
        if (b_k_contig) {
            const int kk = tid & 15;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int nn = (tid >> 4) + 16 * i;
                const int64_t n = n0 + nn, k = k0 + kk;
                Bs[kk][nn] = (n < N && k < K) ? B.load(k, n) : 0.0f;
            }
        } else {
            const int nn = tid & 63;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int kk = (tid >> 6) + 4 * i;
                const int64_t n = n0 + nn, k = k0 + kk;
                Bs[kk][nn] = (n < N && k < K) ? B.load(k, n) : 0.0f;
            }
        }
        __syncthreads();
        const int kmax = (int)((K - k0 < PG_KT) ? (K - k0) : PG_KT);
        for (int kk = 0; kk < kmax; ++kk) {
            const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float a[4] = {av.x, av.y, av.z, av.w};
            const float b[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] += op.pair(a[i], b[j], col[j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t m = m0 + ty * 4 + i;
        if (m >= Mrows) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int64_t n = n0 + tx * 4 + j;
            if (n < N) ep.store(C, ldc, m, n, op.finish(acc[i][j], col[j]));
        }
    }
  }   // tile loop
}

// 1-D grid of at most 8 CTAs per SM (148 SMs assumed for the cap; the kernel loops over its tiles)
inline dim3 pair_gemm_grid(int64_t Mrows, int64_t N)
{
    const int64_t tiles = ((Mrows + PG_BM - 1) / PG_BM) * ((N + PG_BN - 1) / PG_BN);
    return dim3((unsigned)(tiles < 1 ? 1 : (tiles > 148 * 8 ? 148 * 8 : tiles)));
}

}  // namespace b200
