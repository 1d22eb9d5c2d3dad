// conv.cu -- conv form of the approx path:
//   QCustomBNConv2dTorch.run_forward, approx/approx_calculation.py:822-917
//     im2col (:724-747)  ->  per-group approx_multiply (:858-906)  ->  [B*Ho*Wo, C_out]
// The reference materialises the K-times inflated im2col matrix with kh*kw strided copies and then
// loops over groups and output columns in Python.  Here:
//   * groups with N = C_out/groups > 1 : the table GEMM reads the NCHW input through an implicit
//     im2col view in its operand-predecode pass (no im2col matrix in HBM), one GEMM per group;
//   * groups with N == 1 (every depthwise conv): ONE launch for all groups, with the reference's
//     single-column numerics (int32-tensor biases -> integer pow in param_prepare,
//     approx_calculation.py:800-809);
//   * b200_im2col: the explicit matrix, for the exact (non-approx) conv and for tests.
#include "common.cuh"
#include "conv_geom.cuh"
#include "gemm_impl.cuh"
#include "pair_math.cuh"
#include "tables.h"

namespace b200 {

// ---- explicit im2col: [B,C,H,W] -> [B*Ho*Wo, C*kh*kw], K order (c, ky, kx) -----------------------
__global__ void __launch_bounds__(256) im2col_kernel(ASrc a, float* __restrict__ out, int64_t rows, int64_t K)
{
    const int64_t total = rows * K;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const int64_t m = i / K, k = i - m * K;
        out[i] = a.load(m, k);
    }
}

// ---- single-column groups (depthwise) --------------------------------------------------------------
// out[m, g] = sum_{k < Kg, ascending} pair(col[m, g*Kg + k], w[g, k]) with per-group weight bias.
// A 32 (rows) x 32 (groups) tile per CTA: compute with rows across lanes (coalesced input reads along
// wo), transpose through shared memory, store with groups across lanes (coalesced output rows).
struct DwArgs {
    ASrc a;              // c0 = 0; channel of group g is g * Cg + (k / (kh*kw))
    const float* w;      // [groups, Kg]
    float* out;          // [rows, groups]  (ldc = groups = C_out)
    int64_t rows;
    int groups, Kg, Cg;
    BiasSrc bias;
    PairCfg cfg;
    const signed char* T;
};

__global__ void __launch_bounds__(256) approx_dw_kernel(DwArgs d)
{
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t m = (int64_t)blockIdx.x * 32 + tx;
    const FmtParam pA = make_param(d.cfg.E, d.cfg.M, d.bias.A(), d.cfg.quirk);
    const FmtParam pR = make_param(d.cfg.E, d.cfg.M, d.bias.R(), d.cfg.quirk);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gl = ty + 8 * i;
        const int g = blockIdx.y * 32 + gl;
        float acc = 0.0f;
        if (g < d.groups && m < d.rows) {
            const FmtParam pB = make_param(d.cfg.E, d.cfg.M, d.bias.B(g), d.cfg.quirk);
            ASrc a = d.a;
            a.c0 = g * d.Cg;
            const float* wg = d.w + (int64_t)g * d.Kg;
            for (int k = 0; k < d.Kg; ++k) {
                const float av = a.load(m, k);
                const float bv = __ldg(wg + k);
                acc += d.T ? pair_generic(d.cfg, pA, pB, pR, d.T, av, bv)
                           : pair_generic<signed char>(d.cfg, pA, pB, pR, nullptr, av, bv);
            }
        }
        tile[gl][tx] = acc;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ml = ty + 8 * i;
        const int64_t mm = (int64_t)blockIdx.x * 32 + ml;
        const int g = blockIdx.y * 32 + tx;
        if (mm < d.rows && g < d.groups) d.out[mm * d.groups + g] = tile[tx][ml];
    }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_conv_out_shape(const b200_conv_geom* g, int64_t* rows, int32_t* Ho, int32_t* Wo, int32_t* Kg,
                                   int32_t* Ng)
{
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    if (rows) *rows = d.rows;
    if (Ho) *Ho = d.Ho;
    if (Wo) *Wo = d.Wo;
    if (Kg) *Kg = d.Kg;
    if (Ng) *Ng = d.Ng;
    return B200_OK;
}

extern "C" int b200_im2col(const float* x, float* out, const b200_conv_geom* g, void* stream)
{
    if (int rc = check_device()) return rc;
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    const int64_t K = (int64_t)g->C_in * g->kh * g->kw;
    if (d.rows == 0) return B200_OK;
    if (!x || !out) { set_error("b200_im2col: NULL operand"); return B200_ERR_ARGUMENT; }
    ConvDims all = d;
    all.Cg = g->C_in;
    ASrc a = make_conv_asrc(x, g, all, 0);
    int sms = 148;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t blocks = ceil_div(d.rows * K, 256);
    if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
    im2col_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, out, d.rows, K);
    B200_LAUNCH_CHECK("im2col_kernel");
    return B200_OK;
}

extern "C" size_t b200_approx_conv2d_workspace_bytes(const b200_conv_geom* g)
{
    ConvDims d;
    if (conv_dims(g, &d) != B200_OK) return 0;
    if (d.Ng == 1) return 0;
    return b200_approx_gemm_workspace_bytes(d.rows, d.Kg, d.Ng);
}

extern "C" int b200_approx_conv2d_fbias(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                        const float* bA_dev, const float* bB_dev, int64_t bB_stride,
                                        const float* bR_dev, const b200_approx_params* p, void* workspace,
                                        size_t workspace_bytes, void* stream)
{
    if (int rc = check_device()) return rc;
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    if (!p) { set_error("b200_approx_conv2d: params is NULL"); return B200_ERR_ARGUMENT; }
    const int dflt = 1 << (p->expo_width > 0 ? p->expo_width - 1 : 0);   // approx_calculation.py:766-767
    if (d.Ng != 1) {
        b200_approx_params pp = *p;
        pp.single_column_quirk = 0;
        for (int grp = 0; grp < g->groups; ++grp) {
            BiasSrc b{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev ? bB_dev + (int64_t)grp * d.Ng * bB_stride : nullptr,
                      bB_stride};
            const float* wg = w + (int64_t)grp * d.Ng * d.Kg;
            if (int rc = approx_gemm_impl(make_conv_asrc(x, g, d, grp), wg, 1, d.Kg, out + (int64_t)grp * d.Ng,
                                          g->C_out, d.rows, d.Kg, d.Ng, b, &pp, workspace, workspace_bytes, stream))
                return rc;
        }
        return B200_OK;
    }
    // N == 1 per group: the reference's tensor-bias branch (approx_calculation.py:800-809)
    const int E = p->expo_width, M = p->mant_width;
    const signed char* host_table = nullptr;
    if (p->custom_table) {
        if (E < 1 || M < 1 || M > 5 || E + M > 7) { set_error("Invalid combination of expo_width and mant_width"); return B200_ERR_INVALID_FORMAT; }
        host_table = intern_table(reinterpret_cast<const signed char*>(p->custom_table), M);
    } else if (int rc = select_table(E, M, p->withComp, p->dnsmp_factor, p->allow_unpinned_format, &host_table)) {
        return rc;
    }
    if (d.rows == 0) return B200_OK;
    if (!x || !w || !out || !bB_dev) { set_error("b200_approx_conv2d: NULL operand"); return B200_ERR_ARGUMENT; }
    const signed char* dev_table = nullptr;
    if (int rc = device_table(host_table, M, &dev_table)) return rc;
    DwArgs a;
    a.a = make_conv_asrc(x, g, d, 0);
    a.w = w; a.out = out; a.rows = d.rows; a.groups = g->groups; a.Kg = d.Kg; a.Cg = d.Cg;
    a.bias = BiasSrc{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev, bB_stride};
    a.cfg.E = E; a.cfg.M = M;
    a.cfg.with_approx = p->with_approx != 0 && dev_table != nullptr;
    a.cfg.s2nn2s = p->with_s2nn2s_opt != 0;
    a.cfg.clip_OF = p->golden_clip_OF != 0;
    a.cfg.qbma = p->quant_btw_mult_accu != 0;
    a.cfg.quirk = true;
    a.T = a.cfg.with_approx ? dev_table : nullptr;
    const dim3 grid((unsigned)ceil_div(d.rows, 32), (unsigned)ceil_div(g->groups, 32));
    approx_dw_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
    B200_LAUNCH_CHECK("approx_dw_kernel");
    set_path("single-column");
    return B200_OK;
}
