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
#include <algorithm>

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
    Epilogue ep;         // fused +bias -> BN -> activation -> NCHW store (epilogue.cuh); on == 0: plain [rows, groups]
};

__global__ void __launch_bounds__(256) approx_dw_kernel(DwArgs d, const int* __restrict__ gate, int gate_run_if_nonzero)
{
    if (gate && ((*gate != 0) != (gate_run_if_nonzero != 0))) return;
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const FmtParam pA = make_param(d.cfg.E, d.cfg.M, d.bias.A(), d.cfg.quirk);
    const FmtParam pR = make_param(d.cfg.E, d.cfg.M, d.bias.R(), d.cfg.quirk);
    // 1-D grid over the (32 rows x 32 groups) tiles, group tile fastest: the gated launch behind every table-driven
    // depthwise call exits at once and must not cost tens of thousands of CTAs
    const int64_t gtiles = (d.groups + 31) / 32, tiles = ((d.rows + 31) / 32) * gtiles;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t bx = t / gtiles;
    const int by = (int)(t - bx * gtiles);
    const int64_t m = bx * 32 + tx;
    __syncthreads();   // the previous tile's readers of tile[] are done
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gl = ty + 8 * i;
        const int g = by * 32 + gl;
        float acc = 0.0f;
        if (g < d.groups && m < d.rows) {
            const FmtParam pB = make_param(d.cfg.E, d.cfg.M, d.bias.B(g), d.cfg.quirk);
            ASrc a = d.a;
            a.c0 = g * d.Cg;
            const float* wg = d.w + (int64_t)g * d.Kg;
            for (int k = 0; k < d.Kg; ++k) {
                const float av = a.load(m, k);
                const float bv = __ldg(wg + k);
                acc += d.T ? pair_any(d.cfg, pA, pB, pR, d.T, av, bv)
                           : pair_any<signed char>(d.cfg, pA, pB, pR, nullptr, av, bv);
            }
        }
        tile[gl][tx] = acc;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ml = ty + 8 * i;
        const int64_t mm = bx * 32 + ml;
        const int g = by * 32 + tx;
        if (mm < d.rows && g < d.groups) d.ep.store(d.out, d.groups, mm, g, tile[tx][ml]);
    }
  }   // tile loop
}

// ---- single-column groups, table driven -----------------------------------------------------------
// Same function as approx_dw_kernel for the main mode (with_s2nn2s_opt && quant_btw_mult_accu, no
// golden_clip_OF), operands on the FP8 grid and all three exponent biases >= 2 (where the integer pow of
// the tensor-bias branch makes min_norm 0): the addend is u * F(s, t, mA, mB) with the quirk table of
// approx_gemm.cu (signed-index layout, lut.cuh).  Per MAC: field extraction of `a` (the weight word is
// warp-uniform), one 16-bit shared-memory lookup, one fp32 add, k ascending like the arithmetic kernel.
// Anything outside those conditions raises *flag; approx_dw_kernel then recomputes the whole output.
struct DwLutArgs {
    ASrc a;
    const float* w;
    float* out;
    int64_t rows, n_row_tiles;
    int groups, Kg, Cg, n_grp_tiles;
    BiasSrc bias;
    LutGeom g;
    const uint16_t* lut;
    int lut_bytes;
    int* flag;
    Epilogue ep;         // as in DwArgs; the NCHW store needs no transposition here (rows run across lanes)
};

// Weight fields are decoded once per (group, k) of a tile into `wpart` (they are the same for every row), the
// per-k input offsets once per CTA into `kinfo`; the inner loop then only extracts the fields of `a`.
// KG: compile-time Kg (9 = 3x3 depthwise) or 0 for a run-time Kg.  The KG > 0 body is branch free: all KG input
// loads are issued first (memory-level parallelism), then every tap is evaluated under predicates; taps that
// contribute nothing (zero operand, padding) or that leave the table's domain look up an all-zero cell.
constexpr uint32_t kDwSkip = 0xFFFFFFFFu;
template <int KG>
__global__ void __launch_bounds__(256, 3) approx_dw_lut_kernel(DwLutArgs d)
{
    extern __shared__ __align__(16) unsigned char dw_smem[];
    const int Kg = KG ? KG : d.Kg;
    float (*tile)[33] = reinterpret_cast<float (*)[33]>(dw_smem + d.lut_bytes);
    uint32_t* wpart = reinterpret_cast<uint32_t*>(dw_smem + d.lut_bytes + 32 * 33 * sizeof(float));   // [32][Kg]
    int2* kinfo = reinterpret_cast<int2*>(wpart + ((32 * Kg + 1) & ~1));                               // [Kg]
    {
        const uint4* src = reinterpret_cast<const uint4*>(d.lut);
        uint4* dst = reinterpret_cast<uint4*>(dw_smem);
        for (int i = threadIdx.x; i < d.lut_bytes / 16; i += 256) dst[i] = __ldg(src + i);
    }
    const ASrc& s = d.a;
    for (int k = threadIdx.x; k < Kg; k += 256) {   // K order (c, ky, kx), approx_calculation.py:724-747
        const int kx = k % s.kw, t = k / s.kw, ky = t % s.kh, c = t / s.kh;
        kinfo[k] = make_int2((c * s.H + ky * s.dh) * s.W + kx * s.dw, (ky * s.dh) | ((kx * s.dw) << 16));
    }
    const uint16_t* lut_s = reinterpret_cast<const uint16_t*>(dw_smem);
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const LutGeom& g = d.g;
    const int M = g.M;
    const int bA = d.bias.A(), bR = d.bias.R();
    int off = (bA < 2 || bR < 2 || bA >= 64 || bR >= 64) ? 1 : 0;
    const float unit = pow2f_i(1 - bR - M);
    const uint32_t frac_mask = (1u << (23 - M)) - 1u, mmask = (1u << M) - 1u;
    const uint32_t ea_lo = (uint32_t)(127 + 1 - bA - M);   // exponent field of the smallest sub-normal of A's format
    const uint32_t RA = (uint32_t)g.RA, s_stride = (uint32_t)g.s_stride, a_stride = (uint32_t)g.a_stride,
                   row_entries = (uint32_t)g.row_entries, zero_idx = (uint32_t)g.zeroA_off;
    const int mshift = 23 - M;
    const bool small = d.rows < ((int64_t)1 << 31) && d.n_row_tiles * d.n_grp_tiles < ((int64_t)1 << 31);
    int2 kreg[KG > 0 ? KG : 1];
    if (KG > 0) {
        __syncthreads();
#pragma unroll
        for (int k = 0; k < KG; ++k) kreg[k] = kinfo[k];
    }
    for (int64_t t = blockIdx.x; t < d.n_row_tiles * d.n_grp_tiles; t += gridDim.x) {
        int64_t rt;
        int gt;
        if (small) { rt = (uint32_t)t / (uint32_t)d.n_grp_tiles; gt = (int)((uint32_t)t - (uint32_t)rt * (uint32_t)d.n_grp_tiles); }
        else       { rt = t / d.n_grp_tiles; gt = (int)(t - rt * d.n_grp_tiles); }
        __syncthreads();                          // previous tile's wpart / tile readers are done; first pass: lut, kinfo
        for (int i = threadIdx.x; i < 32 * Kg; i += 256) {
            const int gl = i / Kg, k = i - gl * Kg, grp = gt * 32 + gl;
            uint32_t wp = kDwSkip;
            if (grp < d.groups) {
                const int bB = d.bias.B(grp);
                if (bB < 2 || bB >= 64) off = 1;
                const uint32_t wb = __float_as_uint(__ldg(d.w + (int64_t)grp * Kg + k));
                const uint32_t wmag = wb & 0x7FFFFFFFu;
                if (wmag != 0u) {                 // zero weight: addend 0 (v9 :81)
                    const int LB = (int)(wmag >> 23) - 127;
                    const int tB = LB + bR - bA + 3 + g.RA;          // no clamp: every row is live here
                    if ((wmag & frac_mask) || tB < 0 || tB > g.NTB - 1 || LB < 1 - bB - M || (wmag >> 23) == 0 ||
                        (wmag >> 23) == 255)
                        off = 1;                  // outside the table's domain: the arithmetic kernel recomputes
                    else
                        wp = s_stride * (wb >> 31) + row_entries * (uint32_t)tB + ((wmag & 0x7FFFFFu) >> mshift);
                }
            }
            wpart[i] = wp;
        }
        const bool ep_nchw = d.ep.on && d.ep.HW > 0;
        Epilogue::Col* ecol = reinterpret_cast<Epilogue::Col*>(&tile[0][0]);   // the transposition tile is idle then
        if (ep_nchw && threadIdx.x < 32) {
            const int grp = gt * 32 + threadIdx.x;
            ecol[threadIdx.x] = d.ep.column(grp < d.groups ? grp : d.groups - 1);
        }
        __syncthreads();
        const int64_t m = rt * 32 + tx;
        int wo, ho;
        int64_t b;
        if (small) {
            const uint32_t mm = (uint32_t)m, t2 = mm / (uint32_t)s.Wo;
            wo = (int)(mm - t2 * (uint32_t)s.Wo);
            b = t2 / (uint32_t)s.Ho;
            ho = (int)(t2 - (uint32_t)b * (uint32_t)s.Ho);
        } else {
            wo = (int)(m % s.Wo);
            const int64_t t2 = m / s.Wo;
            ho = (int)(t2 % s.Ho);
            b = t2 / s.Ho;
        }
        const int h0 = ho * s.sh - s.ph, w0 = wo * s.sw - s.pw;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int gl = ty + 8 * i;
            const int grp = gt * 32 + gl;
            float acc = 0.0f;
            if (grp < d.groups && m < d.rows) {
                // address of tap (0, 0, 0); may lie outside the image (padding): only in-range taps are read
                const float* p0 = s.p + ((b * s.C + (int64_t)grp * d.Cg) * s.H + h0) * (int64_t)s.W + w0;
                const uint32_t* wrow = wpart + gl * Kg;
                if (KG > 0) {
                    uint32_t ab[KG > 0 ? KG : 1];
#pragma unroll
                    for (int k = 0; k < KG; ++k) {
                        const int h = h0 + (kreg[k].y & 0xFFFF), wi = w0 + (kreg[k].y >> 16);
                        const bool inb = (unsigned)h < (unsigned)s.H && (unsigned)wi < (unsigned)s.W;   // F.pad zeros (:727)
                        ab[k] = inb ? __float_as_uint(__ldg(p0 + kreg[k].x)) : 0u;
                    }
#pragma unroll
                    for (int k = 0; k < KG; ++k) {
                        const uint32_t wp = wrow[k];                             // warp-uniform
                        const uint32_t amag = ab[k] & 0x7FFFFFFFu;
                        const uint32_t tA = (amag >> 23) - ea_lo;                // wraps for fp32 sub-normals
                        const bool live = (wp != kDwSkip) && (amag != 0u);       // else: zero product, addend 0 (v9 :81)
                        const bool bad = live && (((amag & frac_mask) != 0u) || tA > RA);
                        if (bad) off = 1;
                        const uint32_t mA = (amag >> mshift) & mmask;
                        uint32_t idx = wp + s_stride * (ab[k] >> 31) + a_stride * mA + row_entries * tA;
                        idx = (live && !bad) ? idx : zero_idx;
                        acc += __uint_as_float((uint32_t)lut_s[idx] << 16);
                    }
                } else {
                    for (int k = 0; k < Kg; ++k) {
                        const uint32_t wp = wrow[k];                             // warp-uniform
                        if (wp == kDwSkip) continue;
                        const int2 ki = kinfo[k];
                        const int h = h0 + (ki.y & 0xFFFF), wi = w0 + (ki.y >> 16);
                        if ((unsigned)h >= (unsigned)s.H || (unsigned)wi >= (unsigned)s.W) continue;   // F.pad zeros (:727)
                        const uint32_t abk = __float_as_uint(__ldg(p0 + ki.x));
                        const uint32_t amag = abk & 0x7FFFFFFFu;
                        if (amag == 0u) continue;                                // zero product: addend 0 (v9 :81)
                        const uint32_t tA = (amag >> 23) - ea_lo;                // wraps for fp32 sub-normals
                        if ((amag & frac_mask) || tA > RA) { off = 1; continue; }
                        const uint32_t mA = (amag >> mshift) & mmask;
                        const uint32_t idx = wp + s_stride * (abk >> 31) + a_stride * mA + row_entries * tA;
                        acc += __uint_as_float((uint32_t)lut_s[idx] << 16);
                    }
                }
            }
            if (ep_nchw) {
                // rows (= pixels) run across the lanes: the NCHW store is coalesced as it stands
                if (grp < d.groups && m < d.rows)
                    d.out[((b * d.ep.Ctot + d.ep.c0 + grp) * (int64_t)s.Ho + ho) * s.Wo + wo] = d.ep.apply(acc * unit, ecol[gl]);
            } else {
                tile[gl][tx] = acc * unit;
            }
        }
        if (ep_nchw) continue;
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int ml = ty + 8 * i;
            const int64_t mm = rt * 32 + ml;
            const int grp = gt * 32 + tx;
            if (mm < d.rows && grp < d.groups) d.ep.store(d.out, d.groups, mm, grp, tile[tx][ml]);
        }
    }
    if (off) *d.flag = 1;
}

// ---- single-column groups, table driven, over a PRE-DECODED input plane ----------------------------------------
// approx_dw_lut_kernel above extracts and validates the fields of `a` for every MAC (52 instructions per MAC measured,
// profiles/r1e_approx_dw_lut_full.txt) although every input element is used by up to kh*kw outputs.  Here the input is
// decoded ONCE (dw_plane_kernel: optional fused input fake-quantizer, field extraction, validity test -> one 32-bit
// word per element = byte offset of its (sign, tA, mA) part of the table index, same NCHW layout), and the MAC is
//     LDG word -> address = min(word + weight part, zero cell) -> LDS.U16 -> fp32 add          (k ascending)
// A zero / padded / invalid input carries a huge word, which the min() sends to the all-zero cell.
constexpr uint32_t kDwPlaneSkip = 0x40000000u;

struct DwPlaneArgs {
    ASrc a;               // geometry + the raw input (fq fields set: quantize while decoding)
    uint32_t* plane;      // [B][C][H][W] words
    int64_t total;        // B * C * H * W
    BiasSrc bias;
    LutGeom g;
    int* flag;
};

__device__ __forceinline__ uint32_t dw_plane_word(float v, uint32_t frac_mask, uint32_t mmask, uint32_t ea_lo, uint32_t RA,
                                                  uint32_t s_stride2, uint32_t a_stride2, uint32_t row2, int mshift, int& off)
{
    const uint32_t ab = __float_as_uint(v);
    const uint32_t amag = ab & 0x7FFFFFFFu;
    if (amag == 0u) return kDwPlaneSkip;                       // zero product: addend 0 (v9 :81)
    const uint32_t tA = (amag >> 23) - ea_lo;                  // wraps for values below the sub-normal range
    if ((amag & frac_mask) || tA > RA) { off = 1; return kDwPlaneSkip; }
    const uint32_t mA = (amag >> mshift) & mmask;
    return s_stride2 * (ab >> 31) + a_stride2 * mA + row2 * tA;
}

__global__ void __launch_bounds__(256) dw_plane_kernel(DwPlaneArgs d)
{
    const ASrc& s = d.a;
    const LutGeom& g = d.g;
    const int M = g.M, bA = d.bias.A(), bR = d.bias.R();
    int off = (bA < 2 || bR < 2 || bA >= 64 || bR >= 64) ? 1 : 0;
    const uint32_t frac_mask = (1u << (23 - M)) - 1u, mmask = (1u << M) - 1u;
    const uint32_t ea_lo = (uint32_t)(127 + 1 - bA - M), RA = (uint32_t)g.RA;
    const uint32_t s2 = 2u * (uint32_t)g.s_stride, a2 = 2u * (uint32_t)g.a_stride, r2 = 2u * (uint32_t)g.row_entries;
    const int mshift = 23 - M;
    // fused input fake-quantizer: constants read once per thread
    const bool fq = s.fq_maxval != nullptr;
    const float mx = fq ? __ldg(s.fq_maxval) : 0.0f, fb = fq ? __ldg(s.fq_bias) : 0.0f;
    const bool fast = fq && fp8_ste_fast_ok(fb);
    auto quant = [&](float v) -> float {
        if (!fq) return v;
        return fast ? fp8_ste_quant_fast(v, mx, (int)fb, (int)s.fq_Mf, s.fq_sign) : fp8_ste_quant(v, mx, fb, s.fq_Mf, s.fq_sign);
    };
    const bool vec = (d.total & 3) == 0 && ((uintptr_t)s.p & 15) == 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    if (vec) {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < d.total / 4; i += stride) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(s.p) + i);
            uint4 w;
            w.x = dw_plane_word(quant(v.x), frac_mask, mmask, ea_lo, RA, s2, a2, r2, mshift, off);
            w.y = dw_plane_word(quant(v.y), frac_mask, mmask, ea_lo, RA, s2, a2, r2, mshift, off);
            w.z = dw_plane_word(quant(v.z), frac_mask, mmask, ea_lo, RA, s2, a2, r2, mshift, off);
            w.w = dw_plane_word(quant(v.w), frac_mask, mmask, ea_lo, RA, s2, a2, r2, mshift, off);
            reinterpret_cast<uint4*>(d.plane)[i] = w;
        }
    } else {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < d.total; i += stride)
            d.plane[i] = dw_plane_word(quant(__ldg(s.p + i)), frac_mask, mmask, ea_lo, RA, s2, a2, r2, mshift, off);
    }
    if (off) *d.flag = 1;
}

constexpr int kDwPlaneRows = 256;   // output pixels per tile of approx_dw_plane_lut_kernel (one per thread)
struct DwPlaneLutArgs {
    DwLutArgs l;             // a.p is unused here; everything else as in approx_dw_lut_kernel
    const uint32_t* plane;   // [B][C][H][W] words written by dw_plane_kernel
};

// Same tiling as approx_dw_lut_kernel (persistent CTAs over 32 rows x 32 groups, rows across lanes, weight parts decoded
// once per tile into shared memory).  KG = 9: 3x3 taps unrolled, all nine words loaded first.
template <int KG>
__global__ void __launch_bounds__(256, 4) approx_dw_plane_lut_kernel(DwPlaneLutArgs pa)
{
    const DwLutArgs& d = pa.l;
    extern __shared__ __align__(16) unsigned char dw_smem[];
    const int Kg = KG ? KG : d.Kg;
    float (*tile)[33] = reinterpret_cast<float (*)[33]>(dw_smem + d.lut_bytes);
    uint32_t* wpart = reinterpret_cast<uint32_t*>(dw_smem + d.lut_bytes + 32 * 33 * sizeof(float));   // [32][Kg]
    int2* kinfo = reinterpret_cast<int2*>(wpart + ((32 * Kg + 1) & ~1));                               // [Kg]
    {
        const uint4* src = reinterpret_cast<const uint4*>(d.lut);
        uint4* dst = reinterpret_cast<uint4*>(dw_smem);
        for (int i = threadIdx.x; i < d.lut_bytes / 16; i += 256) dst[i] = __ldg(src + i);
    }
    const ASrc& s = d.a;
    for (int k = threadIdx.x; k < Kg; k += 256) {   // K order (c, ky, kx), approx_calculation.py:724-747
        const int kx = k % s.kw, t = k / s.kw, ky = t % s.kh, c = t / s.kh;
        kinfo[k] = make_int2((c * s.H + ky * s.dh) * s.W + kx * s.dw, (ky * s.dh) | ((kx * s.dw) << 16));
    }
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(dw_smem);
    const LutGeom& g = d.g;
    const int M = g.M;
    const int bA = d.bias.A(), bR = d.bias.R();
    int off = (bA < 2 || bR < 2 || bA >= 64 || bR >= 64) ? 1 : 0;
    const float unit = pow2f_i(1 - bR - M);
    const uint32_t frac_mask = (1u << (23 - M)) - 1u;
    const uint32_t zero_addr = lut_base + 2u * (uint32_t)g.zeroA_off;   // an all-zero cell
    const int mshift = 23 - M;
    const bool small = d.rows < ((int64_t)1 << 31) && d.n_row_tiles * d.n_grp_tiles < ((int64_t)1 << 31);
    int2 kreg[KG > 0 ? KG : 1];
    if (KG > 0) {
        __syncthreads();
#pragma unroll
        for (int k = 0; k < KG; ++k) kreg[k] = kinfo[k];
    }
    // Tiles in group-tile-major order: a CTA's consecutive tiles (stride gridDim.x) mostly share their 32 groups, whose
    // decoded weight parts and epilogue constants then stay in shared memory (no decode, no barrier per tile).
    const bool ep_nchw = d.ep.on && d.ep.HW > 0;
    Epilogue::Col* ecol = reinterpret_cast<Epilogue::Col*>(&tile[0][0]);
    int last_gt = -1;
    for (int64_t t = blockIdx.x; t < d.n_row_tiles * d.n_grp_tiles; t += gridDim.x) {
        int64_t rt;
        int gt;
        if (small) { gt = (int)((uint32_t)t / (uint32_t)d.n_row_tiles); rt = (uint32_t)t - (uint32_t)gt * (uint32_t)d.n_row_tiles; }
        else       { gt = (int)(t / d.n_row_tiles); rt = t - (int64_t)gt * d.n_row_tiles; }
        if (gt != last_gt) {
        last_gt = gt;
        __syncthreads();                          // previous tile's wpart / tile readers are done; first pass: lut, kinfo
        for (int i = threadIdx.x; i < 32 * Kg; i += 256) {
            const int gl = i / Kg, k = i - gl * Kg, grp = gt * 32 + gl;
            uint32_t wp = kDwPlaneSkip;           // zero weight (v9 :81) or no such group: every tap looks up the zero cell
            if (grp < d.groups) {
                const int bB = d.bias.B(grp);
                if (bB < 2 || bB >= 64) off = 1;
                const uint32_t wb = __float_as_uint(__ldg(d.w + (int64_t)grp * Kg + k));
                const uint32_t wmag = wb & 0x7FFFFFFFu;
                if (wmag != 0u) {
                    const int LB = (int)(wmag >> 23) - 127;
                    const int tB = LB + bR - bA + 3 + g.RA;          // no clamp: every row is live here
                    if ((wmag & frac_mask) || tB < 0 || tB > g.NTB - 1 || LB < 1 - bB - M || (wmag >> 23) == 0 ||
                        (wmag >> 23) == 255)
                        off = 1;                  // outside the table's domain: the arithmetic kernel recomputes
                    else
                        wp = lut_base + 2u * ((uint32_t)g.s_stride * (wb >> 31) + (uint32_t)g.row_entries * (uint32_t)tB +
                                              ((wmag & 0x7FFFFFu) >> mshift));
                }
            }
            wpart[i] = wp;
        }
        if (ep_nchw && threadIdx.x < 32) {
            const int grp = gt * 32 + threadIdx.x;
            ecol[threadIdx.x] = d.ep.column(grp < d.groups ? grp : d.groups - 1);
        }
        __syncthreads();
        }
        // one thread = one output pixel (consecutive pixels across the 256 threads), all 32 groups of the tile in turn:
        // the pixel's coordinates and tap mask are computed once per 288 MACs
        const int64_t m = rt * kDwPlaneRows + threadIdx.x;
        if (m >= d.rows) continue;
        int wo, ho;
        int64_t b;
        if (small) {
            const uint32_t mm = (uint32_t)m, t2 = mm / (uint32_t)s.Wo;
            wo = (int)(mm - t2 * (uint32_t)s.Wo);
            b = t2 / (uint32_t)s.Ho;
            ho = (int)(t2 - (uint32_t)b * (uint32_t)s.Ho);
        } else {
            wo = (int)(m % s.Wo);
            const int64_t t2 = m / s.Wo;
            ho = (int)(t2 % s.Ho);
            b = t2 / s.Ho;
        }
        const int h0 = ho * s.sh - s.ph, w0 = wo * s.sw - s.pw;
        // tap validity (F.pad zeros, :727) once per output pixel, as a bit mask over k
        uint32_t inb = 0u;
        if (KG > 0) {
#pragma unroll
            for (int k = 0; k < KG; ++k) {
                const int h = h0 + (kreg[k].y & 0xFFFF), wi = w0 + (kreg[k].y >> 16);
                inb |= ((unsigned)h < (unsigned)s.H && (unsigned)wi < (unsigned)s.W) ? (1u << k) : 0u;
            }
        }
        const int ngl = min(32, d.groups - gt * 32);
#pragma unroll 4
        for (int gl = 0; gl < ngl; ++gl) {
            const int grp = gt * 32 + gl;
            float acc = 0.0f;
            const uint32_t* p0 = pa.plane + ((b * s.C + (int64_t)grp * d.Cg) * s.H + h0) * (int64_t)s.W + w0;
            const uint32_t* wrow = wpart + gl * Kg;      // the same address in every lane: broadcast loads
            if (KG > 0) {
                uint32_t aw[KG > 0 ? KG : 1];
#pragma unroll
                for (int k = 0; k < KG; ++k) aw[k] = ((inb >> k) & 1u) ? __ldg(p0 + kreg[k].x) : kDwPlaneSkip;
#pragma unroll
                for (int k = 0; k < KG; ++k) {
                    const uint32_t addr = min(aw[k] + wrow[k], zero_addr);   // skip words overshoot: zero cell
                    unsigned short e;
                    asm("ld.shared.u16 %0, [%1];" : "=h"(e) : "r"(addr));
                    // acc += bf16 cell: one FHFMA (cell * 1.0 is exact, the add rounds once) instead of SHL + FADD
                    asm("fma.rn.f32.bf16 %0, %1, %2, %0;" : "+f"(acc) : "h"(e), "h"((unsigned short)0x3F80));
                }
            } else {
                for (int k = 0; k < Kg; ++k) {
                    const int2 ki = kinfo[k];
                    const int h = h0 + (ki.y & 0xFFFF), wi = w0 + (ki.y >> 16);
                    if ((unsigned)h >= (unsigned)s.H || (unsigned)wi >= (unsigned)s.W) continue;   // F.pad zeros (:727)
                    const uint32_t addr = min(__ldg(p0 + ki.x) + wrow[k], zero_addr);
                    unsigned short e;
                    asm("ld.shared.u16 %0, [%1];" : "=h"(e) : "r"(addr));
                    acc += __uint_as_float((uint32_t)e << 16);
                }
            }
            if (ep_nchw) d.out[((b * d.ep.Ctot + d.ep.c0 + grp) * (int64_t)s.Ho + ho) * s.Wo + wo] = d.ep.apply(acc * unit, ecol[gl]);
            else         d.ep.store(d.out, d.groups, m, grp, acc * unit);
        }
    }
    if (off) *d.flag = 1;
}

// ---- exact grouped conv (non-approx configuration, groups > 1) ---------------------------------------
// out[m, n] = sum_{k < Kg, ascending} col[m, grp(n)*Kg + k] * w[n, k]: the reference's exact branch of the
// group loop (`input_col_group @ weight_group.T`, approx/approx_calculation.py:810-811, :796-797) for every
// group in ONE launch, implicit im2col (no [rows, C*kh*kw] matrix in HBM).  32 rows x 32 output channels per
// CTA: rows across lanes for the gather (coalesced along wo), transposed through shared memory for the store.
__global__ void __launch_bounds__(256) exact_grouped_conv_kernel(ASrc s, const float* __restrict__ w, float* __restrict__ out,
                                                                 int64_t rows, int64_t n_row_tiles, int n_ch_tiles, int C_out,
                                                                 int Kg, int Cg, int Ng)
{
    __shared__ float tile[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int64_t t = blockIdx.x; t < n_row_tiles * n_ch_tiles; t += gridDim.x) {
        const int64_t rt = t / n_ch_tiles;
        const int ct = (int)(t - rt * n_ch_tiles);
        const int64_t m = rt * 32 + tx;
        const int wo = (int)(m % s.Wo);
        const int64_t t2 = m / s.Wo;
        const int ho = (int)(t2 % s.Ho);
        const int64_t b = t2 / s.Ho;
        const int h0 = ho * s.sh - s.ph, w0 = wo * s.sw - s.pw;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int nl = ty + 8 * i;
            const int n = ct * 32 + nl;
            float acc = 0.0f;
            if (n < C_out && m < rows) {
                const int grp = n / Ng;
                const float* wn = w + (int64_t)n * Kg;
                const float* base = s.p + (b * s.C + (int64_t)grp * Cg) * (int64_t)s.H * s.W;
                int kx = 0, ky = 0, c = 0;
                for (int k = 0; k < Kg; ++k) {
                    const int h = h0 + ky * s.dh, wi = w0 + kx * s.dw;
                    float av = 0.0f;
                    if (h >= 0 && h < s.H && wi >= 0 && wi < s.W) av = __ldg(base + ((int64_t)c * s.H + h) * s.W + wi);
                    if (++kx == s.kw) { kx = 0; if (++ky == s.kh) { ky = 0; ++c; } }
                    acc = fmaf(av, __ldg(wn + k), acc);
                }
            }
            tile[nl][tx] = acc;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int ml = ty + 8 * i;
            const int64_t mm = rt * 32 + ml;
            const int n = ct * 32 + tx;
            if (mm < rows && n < C_out) out[mm * C_out + n] = tile[tx][ml];
        }
        __syncthreads();
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

extern "C" int b200_exact_grouped_conv2d(const float* x, const float* w, float* out, const b200_conv_geom* g, void* stream)
{
    if (int rc = check_device()) return rc;
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    if (d.rows == 0) return B200_OK;
    if (!x || !w || !out) { set_error("b200_exact_grouped_conv2d: NULL operand"); return B200_ERR_ARGUMENT; }
    ASrc a = make_conv_asrc(x, g, d, 0);
    int sms = 148;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t n_row_tiles = ceil_div(d.rows, 32);
    const int n_ch_tiles = (int)ceil_div(g->C_out, 32);
    const int64_t blocks = std::min<int64_t>(n_row_tiles * n_ch_tiles, (int64_t)sms * 8);
    exact_grouped_conv_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a, w, out, d.rows, n_row_tiles, n_ch_tiles,
                                                                                   g->C_out, d.Kg, d.Cg, d.Ng);
    B200_LAUNCH_CHECK("exact_grouped_conv_kernel");
    set_path("exact-grouped-conv");
    return B200_OK;
}

extern "C" size_t b200_approx_conv2d_prepared_b_bytes(const b200_conv_geom* g)
{
    ConvDims d;
    if (!g || conv_dims(g, &d) != B200_OK || d.Ng == 1) return 0;
    return (size_t)round_up((int64_t)b200_approx_gemm_prepared_b_bytes(d.Kg, d.Ng), 256) * (size_t)g->groups;
}

extern "C" size_t b200_approx_conv2d_workspace_bytes(const b200_conv_geom* g)
{
    ConvDims d;
    if (conv_dims(g, &d) != B200_OK) return 0;
    // single-column groups: device-side fallback flag + the decoded input plane of approx_dw_plane_lut_kernel
    if (d.Ng == 1) return 256 + (size_t)round_up((int64_t)g->batch * g->C_in * g->H * g->W * 4, 256);
    return b200_approx_gemm_workspace_bytes(d.rows, d.Kg, d.Ng) + conv_gather_workspace_bytes(g->batch, d.Cg, g->H, g->W, d.Kg);
}

static int approx_conv2d_impl(const float* x, const float* w, float* out, const b200_conv_geom* g, const float* bA_dev,
                              const float* bB_dev, int64_t bB_stride, const float* bR_dev, const b200_approx_params* p,
                              const b200_act_quant* aq, void* workspace, size_t workspace_bytes, void* stream,
                              const b200_epilogue* ep = nullptr);

extern "C" int b200_approx_conv2d_fbias_ep(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                           const float* bA_dev, const float* bB_dev, int64_t bB_stride,
                                           const float* bR_dev, const b200_approx_params* p, const b200_act_quant* aq,
                                           const b200_epilogue* ep, void* workspace, size_t workspace_bytes, void* stream)
{
    return approx_conv2d_impl(x, w, out, g, bA_dev, bB_dev, bB_stride, bR_dev, p, aq, workspace, workspace_bytes, stream, ep);
}

extern "C" int b200_approx_conv2d_fbias(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                        const float* bA_dev, const float* bB_dev, int64_t bB_stride,
                                        const float* bR_dev, const b200_approx_params* p, void* workspace,
                                        size_t workspace_bytes, void* stream)
{
    return approx_conv2d_impl(x, w, out, g, bA_dev, bB_dev, bB_stride, bR_dev, p, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int b200_approx_conv2d_fbias_fq(const float* x, const float* w, float* out, const b200_conv_geom* g,
                                           const float* bA_dev, const float* bB_dev, int64_t bB_stride,
                                           const float* bR_dev, const b200_approx_params* p, const b200_act_quant* aq,
                                           void* workspace, size_t workspace_bytes, void* stream)
{
    return approx_conv2d_impl(x, w, out, g, bA_dev, bB_dev, bB_stride, bR_dev, p, aq, workspace, workspace_bytes, stream);
}

static int approx_conv2d_impl(const float* x, const float* w, float* out, const b200_conv_geom* g, const float* bA_dev,
                              const float* bB_dev, int64_t bB_stride, const float* bR_dev, const b200_approx_params* p,
                              const b200_act_quant* aq, void* workspace, size_t workspace_bytes, void* stream,
                              const b200_epilogue* epp)
{
    NvtxRange nvtx_range("b200_approx_conv2d");
    if (int rc = check_device()) return rc;
    ConvDims d;
    if (int rc = conv_dims(g, &d)) return rc;
    if (!p) { set_error("b200_approx_conv2d: params is NULL"); return B200_ERR_ARGUMENT; }
    const int dflt = 1 << (p->expo_width > 0 ? p->expo_width - 1 : 0);   // approx_calculation.py:766-767
    Epilogue ep;
    if (int rc = make_epilogue(epp, &ep)) return rc;
    if (ep.on) {          // conv form: the output is the final NCHW tensor
        ep.HW = (int64_t)d.Ho * d.Wo;
        ep.Ctot = g->C_out;
        if (ep.HW <= 0) return B200_OK;
    }
    if (d.Ng != 1) {
        b200_approx_params pp = *p;
        pp.single_column_quirk = 0;
        const size_t prep_chunk = (size_t)round_up((int64_t)b200_approx_gemm_prepared_b_bytes(d.Kg, d.Ng), 256);
        if (p->prepared_b && p->prepared_b_bytes < prep_chunk * (size_t)g->groups) {
            set_error("b200_approx_conv2d: prepared_b too small (see b200_approx_conv2d_prepared_b_bytes)");
            return B200_ERR_WORKSPACE;
        }
        for (int grp = 0; grp < g->groups; ++grp) {
            if (p->prepared_b) {   // one chunk per group
                pp.prepared_b = reinterpret_cast<char*>(p->prepared_b) + prep_chunk * (size_t)grp;
                pp.prepared_b_bytes = prep_chunk;
            }
            BiasSrc b{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev ? bB_dev + (int64_t)grp * d.Ng * bB_stride : nullptr,
                      bB_stride};
            const float* wg = w + (int64_t)grp * d.Ng * d.Kg;
            ASrc av = make_conv_asrc(x, g, d, grp);
            if (int rc = apply_act_quant(&av, aq)) return rc;
            ep.c0 = grp * d.Ng;
            if (int rc = approx_gemm_impl(av, wg, 1, d.Kg, ep.on ? out : out + (int64_t)grp * d.Ng,
                                          g->C_out, d.rows, d.Kg, d.Ng, b, &pp, workspace, workspace_bytes, stream,
                                          ep.on ? &ep : nullptr))
                return rc;
        }
        return B200_OK;
    }
    // N == 1 per group: the reference's tensor-bias branch (approx_calculation.py:800-809)
    const int64_t in_total = (int64_t)g->batch * g->C_in * g->H * g->W;
    const bool plane_room = workspace && workspace_bytes >= 256 + (size_t)in_total * 4 && ((uintptr_t)workspace & 15) == 0;
    // (a fused input quantizer is honoured by the decoded-plane table kernel and by the arithmetic kernel, whose operand
    // view quantizes on load; the per-MAC decoding table kernel reads the raw tensor and is skipped then)
    const int E = p->expo_width, M = p->mant_width;
    const bool v5 = p->ofuf_mode == 1;      // v5 integer-add semantics (no tensor-bias branch exists in v5)
    const signed char* host_table = nullptr;
    if (p->custom_table) {
        if (E < 1 || M < 1 || M > 5 || E + M > 7) { set_error("Invalid combination of expo_width and mant_width"); return B200_ERR_INVALID_FORMAT; }
        host_table = intern_table(reinterpret_cast<const signed char*>(p->custom_table), M);
    } else if (int rc = v5 ? select_table_v5(E, M, p->withComp, p->dnsmp_factor, &host_table)
                           : select_table(E, M, p->withComp, p->dnsmp_factor, p->allow_unpinned_format, &host_table)) {
        return rc;
    }
    if (d.rows == 0) return B200_OK;
    if (!x || !w || !out || !bB_dev) { set_error("b200_approx_conv2d: NULL operand"); return B200_ERR_ARGUMENT; }
    const signed char* dev_table = nullptr;
    if (int rc = device_table(host_table, M, &dev_table)) return rc;
    DwArgs a;
    a.a = make_conv_asrc(x, g, d, 0);
    if (int rc = apply_act_quant(&a.a, aq)) return rc;     // ASrc::load quantizes: the arithmetic kernel honours it too
    a.w = w; a.out = out; a.rows = d.rows; a.groups = g->groups; a.Kg = d.Kg; a.Cg = d.Cg;
    a.bias = BiasSrc{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev, bB_stride};
    a.cfg.E = E; a.cfg.M = M;
    a.cfg.with_approx = p->with_approx != 0 && dev_table != nullptr;
    a.cfg.s2nn2s = p->with_s2nn2s_opt != 0;
    a.cfg.clip_OF = p->golden_clip_OF != 0;
    a.cfg.qbma = p->quant_btw_mult_accu != 0;
    a.cfg.quirk = true;
    a.T = a.cfg.with_approx ? dev_table : nullptr;
    a.ep = ep;           // c0 = 0: group == output channel
    if (v5) {
        a.cfg.v5 = true; a.cfg.quirk = false;
        a.cfg.sim_ofuf = p->sim_hw_add_OFUF != 0; a.cfg.of_opt = p->with_OF_opt != 0; a.cfg.uf_opt = p->with_UF_opt != 0;
        a.T = dev_table;
    }
    const dim3 grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(ceil_div(d.rows, 32) * ceil_div(g->groups, 32), (int64_t)148 * 8)));
    cudaStream_t st = (cudaStream_t)stream;
    // table path: main mode only; needs 256 B of workspace for the device-side fallback flag
    const bool lut_mode = !v5 && !p->force_generic && p->with_s2nn2s_opt && p->quant_btw_mult_accu && !p->golden_clip_OF &&
                          workspace && workspace_bytes >= 256;
    if (lut_mode) {
        DwLutArgs la;
        int usable = 0;
        if (int rc = get_quirk_lut(E, M, a.cfg.with_approx ? host_table : nullptr, a.T, a.cfg.with_approx, &la.g, &la.lut,
                                   &la.lut_bytes, &usable))
            return rc;
        const size_t smem = (size_t)la.lut_bytes + 32 * 33 * sizeof(float) + (size_t)((32 * d.Kg + 1) & ~1) * 4 + (size_t)d.Kg * 8;
        // the kernel keeps per-k input offsets in 32 bits and tap offsets in 16
        const bool small = (int64_t)d.Cg * g->H * g->W < ((int64_t)1 << 31) && (int64_t)g->kh * g->dil_h < 32768 &&
                           (int64_t)g->kw * g->dil_w < 32768;
        if (usable && small && smem <= 227 * 1024) {
            static PerDeviceOnce once;
            int sms = 0;
            if (int rc = once.get(&sms, []() -> int {
                    B200_CUDA_TRY(cudaFuncSetAttribute(approx_dw_lut_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
                    B200_CUDA_TRY(cudaFuncSetAttribute(approx_dw_lut_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
                    B200_CUDA_TRY(cudaFuncSetAttribute(approx_dw_plane_lut_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
                    B200_CUDA_TRY(cudaFuncSetAttribute(approx_dw_plane_lut_kernel<9>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
                    return B200_OK;
                }))
                return rc;
            la.a = a.a; la.w = w; la.out = out; la.rows = d.rows; la.groups = g->groups; la.Kg = d.Kg; la.Cg = d.Cg;
            la.n_row_tiles = ceil_div(d.rows, 32); la.n_grp_tiles = (int)ceil_div(g->groups, 32);
            la.bias = a.bias; la.flag = reinterpret_cast<int*>(workspace);
            la.ep = ep;
            B200_CUDA_TRY(cudaMemsetAsync(la.flag, 0, 4, st));
            const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(8, (227 * 1024) / (smem + 1024)));
            const int64_t tiles = la.n_row_tiles * la.n_grp_tiles;
            const unsigned blocks = (unsigned)std::min<int64_t>(tiles, (int64_t)sms * per_sm);
            const char* pe = getenv("B200_DW_PLANE");     // experiments / tests: 0 = the per-MAC decoding kernel
            if (plane_room && in_total < ((int64_t)1 << 31) && !(pe && atoi(pe) == 0)) {
                // decode the input once (with the input fake-quantizer, if fused), then the lookup kernel
                DwPlaneArgs dp;
                dp.a = a.a; dp.plane = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(workspace) + 256);
                dp.total = in_total; dp.bias = a.bias; dp.g = la.g; dp.flag = la.flag;
                const unsigned pblocks = (unsigned)std::min<int64_t>(ceil_div(ceil_div(in_total, 4), 256), (int64_t)sms * 16);
                dw_plane_kernel<<<pblocks, 256, 0, st>>>(dp);
                B200_LAUNCH_CHECK("dw_plane_kernel");
                DwPlaneLutArgs pl{la, dp.plane};
                pl.l.n_row_tiles = ceil_div(d.rows, kDwPlaneRows);
                const unsigned pblk = (unsigned)std::min<int64_t>(pl.l.n_row_tiles * la.n_grp_tiles, (int64_t)sms * per_sm);
                if (d.Kg == 9) approx_dw_plane_lut_kernel<9><<<pblk, 256, smem, st>>>(pl);
                else           approx_dw_plane_lut_kernel<0><<<pblk, 256, smem, st>>>(pl);
                B200_LAUNCH_CHECK("approx_dw_plane_lut_kernel");
                approx_dw_kernel<<<grid, 256, 0, st>>>(a, la.flag, 1);
                B200_LAUNCH_CHECK("approx_dw_kernel(conditional)");
                set_path("single-column-lut-plane");
                return B200_OK;
            }
            if (!aq) {
                if (d.Kg == 9) approx_dw_lut_kernel<9><<<blocks, 256, smem, st>>>(la);
                else           approx_dw_lut_kernel<0><<<blocks, 256, smem, st>>>(la);
                B200_LAUNCH_CHECK("approx_dw_lut_kernel");
                // runs only if the table kernel met an operand / bias outside its domain
                approx_dw_kernel<<<grid, 256, 0, st>>>(a, la.flag, 1);
                B200_LAUNCH_CHECK("approx_dw_kernel(conditional)");
                set_path("single-column-lut");
                return B200_OK;
            }
        }
    }
    approx_dw_kernel<<<grid, 256, 0, st>>>(a, nullptr, 0);
    B200_LAUNCH_CHECK("approx_dw_kernel");
    set_path(v5 ? "single-column-v5" : "single-column");
    return B200_OK;
}
