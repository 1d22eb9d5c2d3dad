// elementwise.cu -- HBM-bound elementwise kernels of the path:
//   K4  FP8 fake-quantizer   quantize_to_fp8_ste_MM, quantization/quantizers/fp8_quantizer.py:97-173
//   K5  field codec          approx/approx_matmul_whole_v9.py:233-362, approx/fp_any_int_v5.py:55-162
// 128-bit vectorised loads/stores, grid = multiple of the SM count, grid-stride loops.
// Algorithmic traffic: 8 B per element (4 read + 4 written).
#include "common.cuh"
#include "pair_math.cuh"

namespace b200 {

static int g_sm_count = 0;

static int elementwise_grid(int64_t work_items, int threads)
{
    if (g_sm_count == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    int64_t blocks = ceil_div(work_items, threads);
    int64_t cap = (int64_t)g_sm_count * 8;   // 8 resident CTAs of 256 threads per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ---- K4 -------------------------------------------------------------------------------------
__global__ void fp8_bias_kernel(const float* __restrict__ maxval, int64_t n_channels, int n_bits,
                                float mantissa_bits, int sign_bits, float* __restrict__ bias_out)
{
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c < n_channels) bias_out[c] = fp8_ste_bias(maxval[c], n_bits, mantissa_bits, sign_bits, nullptr);
}

// per-tensor: every thread derives the (same) bias once; vectorised body.
__global__ void __launch_bounds__(256)
fp8_fakequant_tensor_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n,
                            const float* __restrict__ maxval, int n_bits, float mantissa_bits,
                            int sign_bits, float* __restrict__ bias_out, int vec_ok)
{
    float Mf;
    const float mx = __ldg(maxval);
    const float bias = fp8_ste_bias(mx, n_bits, mantissa_bits, sign_bits, &Mf);
    if (blockIdx.x == 0 && threadIdx.x == 0) bias_out[0] = bias;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const bool fast = fp8_ste_fast_ok(bias);          // exponent-field form: no log2f, no division (pair_math.cuh)
    const int bias_i = (int)bias, M_i = (int)Mf;
    auto q = [&](float v) { return fast ? fp8_ste_quant_fast(v, mx, bias_i, M_i, sign_bits) : fp8_ste_quant(v, mx, bias, Mf, sign_bits); };
    if (vec_ok) {
        const int64_t n4 = n >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        float4* y4 = reinterpret_cast<float4*>(y);
        for (int64_t i = tid; i < n4; i += stride) {
            float4 v = __ldcs(x4 + i);
            v.x = q(v.x);
            v.y = q(v.y);
            v.z = q(v.z);
            v.w = q(v.w);
            y4[i] = v;
        }
        for (int64_t i = (n4 << 2) + tid; i < n; i += stride) y[i] = q(x[i]);
    } else {
        for (int64_t i = tid; i < n; i += stride) y[i] = q(x[i]);
    }
}

// per-channel (maxval viewed as [C,1,..,1], fp8_quantizer.py:114-115): channel = (i / inner) % C
__global__ void __launch_bounds__(256)
fp8_fakequant_channel_kernel(const float* __restrict__ x, float* __restrict__ y, int64_t n,
                             const float* __restrict__ maxval, const float* __restrict__ bias,
                             int64_t n_channels, int64_t inner, int n_bits, float mantissa_bits,
                             int sign_bits, int vec_ok)
{
    float Mf = rintf(mantissa_bits);
    Mf = fminf(fmaxf(Mf, 1.0f), (float)(n_bits - sign_bits));
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const bool small = n < ((int64_t)1 << 32) && inner < ((int64_t)1 << 32) && n_channels < ((int64_t)1 << 32);
    if (vec_ok) {   // inner % 4 == 0: a float4 never straddles two channels
        const int64_t n4 = n >> 2;
        const float4* x4 = reinterpret_cast<const float4*>(x);
        float4* y4 = reinterpret_cast<float4*>(y);
        for (int64_t i = tid; i < n4; i += stride) {
            // (32-bit division where the tensor allows it: a 64-bit one costs more than the quantizer itself)
            const int64_t c = small ? (int64_t)((((uint32_t)i << 2) / (uint32_t)inner) % (uint32_t)n_channels)
                                    : ((i << 2) / inner) % n_channels;
            const float mx = __ldg(maxval + c), b = __ldg(bias + c);
            float4 v = __ldcs(x4 + i);
            if (fp8_ste_fast_ok(b)) {      // exponent-field form: no log2f, no division (pair_math.cuh)
                const int bi = (int)b, Mi = (int)Mf;
                v.x = fp8_ste_quant_fast(v.x, mx, bi, Mi, sign_bits);
                v.y = fp8_ste_quant_fast(v.y, mx, bi, Mi, sign_bits);
                v.z = fp8_ste_quant_fast(v.z, mx, bi, Mi, sign_bits);
                v.w = fp8_ste_quant_fast(v.w, mx, bi, Mi, sign_bits);
            } else {
                v.x = fp8_ste_quant(v.x, mx, b, Mf, sign_bits);
                v.y = fp8_ste_quant(v.y, mx, b, Mf, sign_bits);
                v.z = fp8_ste_quant(v.z, mx, b, Mf, sign_bits);
                v.w = fp8_ste_quant(v.w, mx, b, Mf, sign_bits);
            }
            y4[i] = v;
        }
    } else {
        for (int64_t i = tid; i < n; i += stride) {
            const int64_t c = small ? (int64_t)(((uint32_t)i / (uint32_t)inner) % (uint32_t)n_channels) : (i / inner) % n_channels;
            y[i] = fp8_ste_quant(x[i], __ldg(maxval + c), __ldg(bias + c), Mf, sign_bits);
        }
    }
}

// ---- K5 -------------------------------------------------------------------------------------
__global__ void fpany_encode_kernel(const float* __restrict__ v, int64_t n, int E, int M, int bias, int mode,
                                    int quirk, int clip, int* __restrict__ expo, int* __restrict__ mant)
{
    const FmtParam p = make_param(E, M, bias, quirk != 0);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        int e, m;
        if (mode == 0) fpany_encode(p, v[i], clip != 0, e, m);
        else           allnorm_encode(E, M, bias, v[i], clip != 0, e, m);
        expo[i] = e;
        mant[i] = m;
    }
}

__global__ void fpany_decode_kernel(const int* __restrict__ expo, const int* __restrict__ mant,
                                    const float* __restrict__ sign, int64_t n, int E, int M, int bias, int mode,
                                    float* __restrict__ out)
{
    const FmtParam p = make_param(E, M, bias, false);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float s = sign ? sign[i] : 1.0f;
        out[i] = (mode == 0) ? fpany_decode(p, s, expo[i], mant[i]) : allnorm_decode(M, bias, s, expo[i], mant[i]);
    }
}

__global__ void fpany_quant_kernel(const float* __restrict__ v, int64_t n, int E, int M, int bias, int quirk,
                                   int clip, float* __restrict__ out)
{
    const FmtParam p = make_param(E, M, bias, quirk != 0);
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = fpany_quant(p, v[i], clip != 0);
}

// approx_mult_new (approx/approx_matmul_whole_v5.py:155-182): integer-add approximate multiply on the
// code domain with the simulated hardware adder overflow / underflow.  x_int = (expo << M) + mant.
__global__ void v5_ofuf_kernel(const int32_t* __restrict__ x_int, const int32_t* __restrict__ y_int,
                               const int32_t* __restrict__ x_mant, const int32_t* __restrict__ y_mant,
                               const signed char* __restrict__ table, int64_t n, int B_neg, int E, int M, int sim,
                               int of_opt, int uf_opt, int32_t* __restrict__ out)
{
    const int32_t mod = 1 << (E + M), max_norm_int = mod - 1, mant_scale = 1 << M, tn = 1 << M;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        int mx = x_mant ? x_mant[i] : (x_int[i] & (mant_scale - 1));
        int my = y_mant ? y_mant[i] : (y_int[i] & (mant_scale - 1));
        if (mx < 0) mx += tn;                                    // Python negative indexing wraps once
        if (my < 0) my += tn;
        const int comp = (table && mx >= 0 && mx < tn && my >= 0 && my < tn) ? (int)table[mx * tn + my] : 0;
        int32_t t = x_int[i] + y_int[i] + B_neg + comp;          // :160
        if (sim) {
            const bool of = t > max_norm_int, uf = t < 0;        // :166-167
            t = ((t % mod) + mod) % mod;                         // torch `%` is a floor-mod, :170
            if (of_opt && of) t = max_norm_int;                  // :172-173
            if (uf_opt && uf) t = t % mant_scale;                // :175-176
        }
        out[i] = t;
    }
}

// ---- eval-mode BatchNorm + activation epilogue (quantization/quantized_folded_bn.py:66-78) -----------
// y = act(fma(x, alpha, beta)),  invstd = 1 / sqrt(var + eps),  alpha = gamma * invstd,
// beta' = fma(-mean, alpha, beta): op for op what ATen's CPU batch_norm evaluates in inference mode
// (checked bit for bit against torch 2.11 on this container's CPU, tests/golden/make_bn.py), so that a
// model forward stays bit-identical to the reference run on CPU tensors; cuDNN's kernel rounds
// differently in ~15 % of the elements.  Input: any 4-D strided view (the hot path produces channels-last
// [B*Ho*Wo, C] tensors viewed as NCHW); output: contiguous NCHW -- the layout change that the next
// layer would otherwise pay for with a copy is done here, through a 32 x 32 shared-memory tile.
struct BnArgs {
    const float* x;
    int64_t sb, sc, sp;     // element strides of x: batch, channel, pixel (h*W + w must be linear: sh == W * sw)
    float* y;               // [B, C, HW] contiguous
    int64_t B;
    int C;
    int64_t HW;
    const float *mean, *var, *gamma, *beta;
    float eps;
    int act;                // 0: none, 1: ReLU, 2: ReLU6
};

__device__ __forceinline__ float bn_act_apply(float v, float alpha, float beta, int act)
{
    float r = fmaf(v, alpha, beta);
    if (act >= 1) r = fmaxf(r, 0.0f);
    if (act == 2) r = fminf(r, 6.0f);
    return r;
}

// grid: (ceil(HW/32), ceil(C/32), B); block 256 = 32 x 8.  TRANSPOSE: x is channel-fastest.
template <bool TRANSPOSE>
__global__ void __launch_bounds__(256) bn_act_kernel(BnArgs a)
{
    __shared__ float tile[32][33];
    __shared__ float s_alpha[32], s_beta[32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t p0 = (int64_t)blockIdx.x * 32;
    const int c0 = blockIdx.y * 32;
    const int64_t b = blockIdx.z;
    if (threadIdx.x < 32) {
        const int c = c0 + threadIdx.x;
        float alpha = 0.0f, beta = 0.0f;
        if (c < a.C) {
            const float invstd = 1.0f / sqrtf(__ldg(a.var + c) + a.eps);
            alpha = (a.gamma ? __ldg(a.gamma + c) : 1.0f) * invstd;
            beta = fmaf(-__ldg(a.mean + c), alpha, a.beta ? __ldg(a.beta + c) : 0.0f);
        }
        s_alpha[threadIdx.x] = alpha;
        s_beta[threadIdx.x] = beta;
    }
    __syncthreads();
    const float* xb = a.x + b * a.sb;
    float* yb = a.y + b * (int64_t)a.C * a.HW;
    if (TRANSPOSE) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {            // read: lanes along channels
            const int64_t p = p0 + ty + i * 8;
            const int c = c0 + tx;
            float v = 0.0f;
            if (p < a.HW && c < a.C) v = bn_act_apply(__ldg(xb + c * a.sc + p * a.sp), s_alpha[tx], s_beta[tx], a.act);
            tile[ty + i * 8][tx] = v;            // [pixel][channel]
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {            // write: lanes along pixels
            const int c = c0 + ty + i * 8;
            const int64_t p = p0 + tx;
            if (p < a.HW && c < a.C) yb[(int64_t)c * a.HW + p] = tile[tx][ty + i * 8];
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {            // pixel-fastest input: lanes along pixels for both
            const int cl = ty + i * 8, c = c0 + cl;
            const int64_t p = p0 + tx;
            if (p < a.HW && c < a.C)
                yb[(int64_t)c * a.HW + p] = bn_act_apply(__ldg(xb + c * a.sc + p * a.sp), s_alpha[cl], s_beta[cl], a.act);
        }
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_bn_act(const float* x, int64_t sb, int64_t sc, int64_t sp, float* y, int64_t B, int C, int64_t HW,
                           const float* mean, const float* var, const float* gamma, const float* beta, float eps, int act,
                           void* stream)
{
    NvtxRange nvtx_range("b200_bn_act");
    if (int rc = check_device()) return rc;
    if (B < 0 || C < 1 || HW < 0 || act < 0 || act > 2) { set_error("b200_bn_act: bad argument"); return B200_ERR_ARGUMENT; }
    if (B == 0 || HW == 0) return B200_OK;
    if (!x || !y || !mean || !var) { set_error("b200_bn_act: NULL operand"); return B200_ERR_ARGUMENT; }
    if (B > 65535) { set_error("b200_bn_act: batch too large for one launch"); return B200_ERR_ARGUMENT; }
    BnArgs a{x, sb, sc, sp, y, B, C, HW, mean, var, gamma, beta, eps, act};
    const dim3 grid((unsigned)ceil_div(HW, 32), (unsigned)ceil_div(C, 32), (unsigned)B);
    if (sc == 1 && sp != 1) bn_act_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    else                    bn_act_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    B200_LAUNCH_CHECK("bn_act_kernel");
    return B200_OK;
}

extern "C" int b200_approx_mult_v5(const int32_t* x_int, const int32_t* y_int, const int32_t* x_mant,
                                   const int32_t* y_mant, const int8_t* table_dev, int64_t n, int B_neg, int E, int M,
                                   int sim_hw_add_OFUF, int with_OF_opt, int with_UF_opt, int32_t* out, void* stream)
{
    if (int rc = check_device()) return rc;
    if (E < 1 || M < 1 || E + M > 30) { set_error("Invalid combination of expo_width and mant_width"); return B200_ERR_INVALID_FORMAT; }
    if (n <= 0) return B200_OK;
    if (!x_int || !y_int || !out) { set_error("b200_approx_mult_v5: NULL operand"); return B200_ERR_ARGUMENT; }
    v5_ofuf_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(
        x_int, y_int, x_mant, y_mant, reinterpret_cast<const signed char*>(table_dev), n, B_neg, E, M, sim_hw_add_OFUF,
        with_OF_opt, with_UF_opt, out);
    B200_LAUNCH_CHECK("v5_ofuf_kernel");
    return B200_OK;
}

extern "C" int b200_fp8_fakequant(const float* x, float* y, int64_t n, const float* maxval, int64_t n_channels,
                                  int64_t inner, int n_bits, float mantissa_bits, int sign_bits, float* bias_out,
                                  void* stream)
{
    NvtxRange nvtx_range("b200_fp8_fakequant");
    if (int rc = check_device()) return rc;
    if (n < 0 || n_channels < 1 || !maxval || !bias_out || (n > 0 && (!x || !y))) {
        set_error("b200_fp8_fakequant: bad argument");
        return B200_ERR_ARGUMENT;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const bool aligned = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
    if (n_channels == 1) {
        const int grid = elementwise_grid(ceil_div(n > 0 ? n : 1, 4), 256);
        fp8_fakequant_tensor_kernel<<<grid, 256, 0, st>>>(x, y, n, maxval, n_bits, mantissa_bits, sign_bits,
                                                         bias_out, aligned ? 1 : 0);
        B200_LAUNCH_CHECK("fp8_fakequant_tensor_kernel");
    } else {
        if (inner < 1) {
            set_error("b200_fp8_fakequant: inner must be >= 1 for per-channel maxval");
            return B200_ERR_ARGUMENT;
        }
        fp8_bias_kernel<<<(int)ceil_div(n_channels, 128), 128, 0, st>>>(maxval, n_channels, n_bits, mantissa_bits,
                                                                       sign_bits, bias_out);
        B200_LAUNCH_CHECK("fp8_bias_kernel");
        if (n > 0) {
            const int vec = (aligned && (inner % 4 == 0) && (n % 4 == 0)) ? 1 : 0;
            const int grid = elementwise_grid(vec ? n / 4 : n, 256);
            fp8_fakequant_channel_kernel<<<grid, 256, 0, st>>>(x, y, n, maxval, bias_out, n_channels, inner, n_bits,
                                                              mantissa_bits, sign_bits, vec);
            B200_LAUNCH_CHECK("fp8_fakequant_channel_kernel");
        }
    }
    return B200_OK;
}

extern "C" int b200_fpany_encode(const float* v, int64_t n, int E, int M, int bias, int mode, int quirk, int clip,
                                 int32_t* expo_out, int32_t* mant_out, void* stream)
{
    if (int rc = check_device()) return rc;
    if (n <= 0) return B200_OK;
    fpany_encode_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(v, n, E, M, bias, mode, quirk,
                                                                                  clip, expo_out, mant_out);
    B200_LAUNCH_CHECK("fpany_encode_kernel");
    return B200_OK;
}

extern "C" int b200_fpany_decode(const int32_t* expo, const int32_t* mant, const float* sign, int64_t n, int E, int M,
                                 int bias, int mode, float* out, void* stream)
{
    if (int rc = check_device()) return rc;
    if (n <= 0) return B200_OK;
    fpany_decode_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(expo, mant, sign, n, E, M, bias,
                                                                                  mode, out);
    B200_LAUNCH_CHECK("fpany_decode_kernel");
    return B200_OK;
}

extern "C" int b200_fpany_quant(const float* v, int64_t n, int E, int M, int bias, int quirk, int clip, float* out,
                                void* stream)
{
    if (int rc = check_device()) return rc;
    if (n <= 0) return B200_OK;
    fpany_quant_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(v, n, E, M, bias, quirk, clip, out);
    B200_LAUNCH_CHECK("fpany_quant_kernel");
    return B200_OK;
}
