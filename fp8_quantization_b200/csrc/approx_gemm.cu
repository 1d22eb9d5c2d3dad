// approx_gemm.cu -- K1: the approx_v9 simulated-FP8 GEMM.
//
// Replaces the per-column Python loop approx/approx_calculation.py:946-971 (Linear) / :774-799
// (BN-conv) over custom_matmul_vectorize (approx/approx_matmul_whole_v9.py:10-169).
//
// Two kernels implement the same function:
//   * approx_gemm_generic_kernel : evaluates pair_generic() (pair_math.cuh) per operand pair; valid
//     for ANY fp32 operands and every flag combination, including the single-column tensor-bias
//     quirk.  Slow (tens of fp32/int ops per pair); it is the fallback and the definition.
//   * approx_gemm_lut_kernel     : table-driven register-blocked CUDA-core tile for the main mode
//     (with_s2nn2s_opt && quant_btw_mult_accu) when every operand lies on the FP8 grid, which is
//     what the fake-quantizer guarantees (lut.cuh).  Per pair: IADD, LDS.U16, shift, FFMA.
// The choice between them is made ON THE DEVICE (no host sync): the predecode kernels raise a flag
// if any operand is off-grid; both GEMM kernels are enqueued and the wrong one exits at once.
//
// Accumulation: fp32, k ascending, one thread owns an output element for the whole K loop -- the
// same order as the oracle, hence bit-identical to it; identical to the reference whenever fp32
// summation is exact (always, in the calibrated quant_btw_mult_accu regime: DESIGN.md).
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "common.cuh"
#include "lut.cuh"
#include "pair_gemm.cuh"
#include "pair_math.cuh"
#include "qamaa_op.cuh"
#include "tables.h"
#include "epilogue.cuh"
#include "gemm_impl.cuh"
#include "lut_gemm.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// Pair-table builder: one thread per table cell, evaluating the reference op chain in "unit" space
// (value / u, i.e. with result bias 1-M so that u == 1).
// ------------------------------------------------------------------------------------------------
struct LutFlags {
    int sign_quirk;        // cells where F is not odd in the operand sign (v9 :68 after :36)
    int unrepresentable;   // cells whose value does not fit bf16 exactly
};

// quirk = 1: the single-column (int32-tensor-bias) numerics of approx_calculation.py:800-809 for biases >= 2:
// integer pow makes min_norm 0 in every format (approx_matmul_whole_v9.py:202), so nothing is treated
// as sub-normal, no product rounds to zero (all rows are live) and an exponent code of exactly 0 is
// DECODED as a sub-normal (:321-323).
__global__ void lut_build_kernel(LutGeom g, const signed char* __restrict__ T, int has_table, int with_approx,
                                 int clip_OF, int quirk, uint16_t* __restrict__ lut, LutFlags* __restrict__ flags)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= g.total_entries) return;
    uint16_t out = 0;
    if (idx < g.zeroA_off) {
        const int M = g.M;
        const int mB = idx & ((1 << M) - 1);
        const int row = idx >> M;
        const int idx_t = row % g.NT;
        const int q = row / g.NT;
        const int mA = q & ((1 << M) - 1);
        const int s0 = (q >> M) & 1;   // plain layout: [a<0]; signed-index layout: region & 1 = [a<0] xor [b<0]
        const int t = idx_t - g.RA - M - 2;
        if (idx_t > g.RA || quirk) {
            FmtParam pR = make_param(g.E, M, 1 - M, false);   // unit space: u = 2^(1-bias-M) = 1
            if (quirk) pR.min_norm = 0.0f;
            const float s = ldexpf(1.0f, -M);
            const float pex = (1.0f + (float)mA * s) * (1.0f + (float)mB * s);
            const float p = has_table ? mult_mant(M, mA, mB, T, with_approx != 0)
                                      : mult_mant<signed char>(M, mA, mB, nullptr, false);
            const float scale = pow2f_i(t + M - 1);
            float val[2];
#pragma unroll
            for (int sb = 0; sb < 2; ++sb) {
                const float sg = ((s0 ^ sb) & 1) ? -1.0f : 1.0f;
                float gq = pex * scale * sg;                       // golden product, exact
                gq = fpany_quant(pR, gq, clip_OF != 0);            // v9 :36
                const float sgn = (gq < 0.0f) ? -1.0f : 1.0f;      // v9 :68
                float r = scale * p * sgn;                         // v9 :73-77 (+ :79-80 folded in t)
                val[sb] = fpany_quant(pR, r, clip_OF != 0);        // v9 :108
            }
            if (!g.sidx && val[1] != -val[0]) atomicAdd(&flags->sign_quirk, 1);   // not odd in the sign of b
            const uint32_t bits = __float_as_uint(val[0]);
            if (bits & 0xFFFFu) atomicAdd(&flags->unrepresentable, 1);
            out = (uint16_t)(bits >> 16);
        }
    }
    lut[idx] = out;
}

// Pair table of the v5 integer-add GEMM (ofuf_mode 1; lut.cuh make_lut_geom_v5): the addend is
// sign(a) * sign(b) * dec(t) with t = (eA + eB - c) * 2^M + mA + mB + comp[mA, mB] followed by the adder's
// overflow / underflow simulation (approx_matmul_whole_v5.py:155-182) -- a function of (eA + eB - c, mA, mB), i.e.
// the same additive index structure as the approx_v9 table with the CODE exponents in place of floor(log2).
// Cells are in units of u (result bias 1 - M makes u == 1), bf16; every cell has <= M + 1 significant bits.
__global__ void lut_build_v5_kernel(LutGeom g, const signed char* __restrict__ T, int sim, int of_opt, int uf_opt,
                                    uint16_t* __restrict__ lut, LutFlags* __restrict__ flags)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= g.total_entries) return;
    uint16_t out = 0;
    if (idx < g.zeroA_off) {
        const int M = g.M, ms = 1 << M, mod = 1 << (g.E + M);
        const int mB = idx & (ms - 1);
        const int row = idx >> M;
        const int r = row % g.NT;
        const int q = row / g.NT;
        const int mA = q & (ms - 1);
        const int sa = (q >> M) & 1;
        if (r > g.RA) {
            const int esum = r - (g.v5_lo + g.RA + 1);               // eA + eB - (bA + bB - bR)
            int t = esum * ms + mA + mB + (T ? (int)T[mA * ms + mB] : 0);
            if (sim) {
                const bool of = t > mod - 1, uf = t < 0;
                t = ((t % mod) + mod) % mod;
                if (of_opt && of) t = mod - 1;
                if (uf_opt && uf) t = t % ms;
            }
            const FmtParam pR = make_param(g.E, M, 1 - M, false);   // unit space
            const float v = fpany_decode(pR, sa ? -1.0f : 1.0f, t >> M, t & (ms - 1));
            const uint32_t bits = __float_as_uint(v);
            if ((bits & 0xFFFFu) || !(fabsf(v) < INFINITY)) atomicAdd(&flags->unrepresentable, 1);
            out = (uint16_t)(bits >> 16);
        }
    }
    lut[idx] = out;
}

// Pair table of the quantize_after_mult_and_add mode (approx_calculation.py:959-967): the addend is
// Q(a * b) with Q = quantize_to_fp8_ste_MM(., n_bits, maxval, mantissa_bits, sign_bits); for operands on
// the FP8 grid that is sb * u * F(sa, t, mA, mB) with the SAME index geometry as the approx table
// (t = LA + LB + bias_R, u = 2^(1 - bias_R - M)), where bias_R follows from the result quantizer's maxval.
// maxval lives on the device and moves during calibration, so the table is rebuilt per call (<= 60 K cells,
// a few microseconds) into the caller's workspace; the kernel also publishes bias_R (as fp32) for the
// operand predecode, and raises *flag when the table path does not apply (unsigned grid, exotic widths).
__global__ void lut_build_qamaa_kernel(LutGeom g, QamaaOp op, uint16_t* __restrict__ lut, float* __restrict__ bias_out,
                                       int* __restrict__ flag)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    op.begin();
    if (idx == 0) {
        *bias_out = op.bias;
        if (!op.fast || op.M_i != g.M || op.sign_bits != 1) *flag = 1;
    }
    if (idx >= g.total_entries) return;
    uint16_t out = 0;
    if (idx < g.zeroA_off && op.fast) {
        const int M = g.M;
        const int mB = idx & ((1 << M) - 1);
        const int row = idx >> M;
        const int idx_t = row % g.NT;
        const int q = row / g.NT;
        const int mA = q & ((1 << M) - 1);
        const int sa = (q >> M) & 1;
        const int t = idx_t - g.RA - M - 2;                  // LA + LB + bias_R
        if (idx_t > g.RA) {
            const float s = ldexpf(1.0f, -M);
            const float p = (1.0f + (float)mA * s) * (1.0f + (float)mB * s);     // exact
            const float x = p * pow2f_i(t - op.bias_i) * (sa ? -1.0f : 1.0f);   // a * b for b > 0
            const float v = op.quant(x) * pow2f_i(M + op.bias_i - 1);           // in units of u
            const uint32_t bits = __float_as_uint(v);
            if ((bits & 0xFFFFu) || op.quant(-x) != -op.quant(x)) *flag = 1;    // not bf16-exact / not odd in b
            out = (uint16_t)(bits >> 16);
        }
    }
    lut[idx] = out;
}

// Byte-cell companion of a pair table (lut_gemm.cuh, C8): cell8 = the 16-bit cell where it is an integer of
// magnitude <= 127 (units of u), else 0, and *bad = NT - (first table row holding a cell that does not fit).
// The table GEMM uses byte cells only for rows below that one.  `flag16` (qamaa: the 16-bit table was declared
// unusable on the device) is passed through unchanged.
__global__ void lut8_build_kernel(LutGeom g, const uint16_t* __restrict__ lut16, int8_t* __restrict__ lut8, int n8,
                                  int* __restrict__ bad)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n8) return;
    int8_t out = 0;
    if (idx < g.zeroA_off) {
        const float v = __uint_as_float((uint32_t)lut16[idx] << 16);
        if (v == rintf(v) && fabsf(v) <= 127.0f) out = (int8_t)(int)v;
        else atomicMax(bad, g.NT - (idx >> g.M) % g.NT);
    }
    lut8[idx] = out;
}

constexpr int kEscLists = 32;

// Correction pass of the byte-cell variant: for every listed escape A element (row m, k) and every column n, the
// pair's 16-bit cell is looked up in the full table; where the byte table holds 0 in its place (the cell is not an
// integer within +-127) the cell, in units of u, is added to D[m][n].  The byte-cell GEMM starts its sums from D.
// Work is (escapes x N) cheap tests and rare lookups: escapes are at most 1/8 of the A elements (usually a 1e-4
// fraction), or the gate is closed and nobody reads D.  Block 0 also publishes the total and the overflow flag.
// Two-sided escapes.  A pair needs a 16-bit cell only if tA_idx + tB_idx lies above the last all-int8 row, so what
// counts as an A escape depends on the largest tB_idx -- and one thinly populated top binade of B (the channel maxima)
// can turn a whole populated binade of A into escapes.  c8_pick_kernel therefore reads the histogram of tB_idx and
// declares the top levels holding at most 1/256 of B "rare"; c8_brare_kernel lists those elements; the A pass compares
// against the largest NON-rare level, and the correction pass covers (A escape x any B) and (rare B x non-escape A).
struct C8CorrectArgs {
    const unsigned long long* esc_list;
    unsigned esc_cap;
    const unsigned* esc_counters;
    unsigned* esc_total;       // [0] total, [1] a list overflowed
    const unsigned long long* brare_list;
    unsigned brare_cap;
    const unsigned* brare_count;
    const uint32_t* Aw;
    const uint32_t* Bw;
    int64_t Mpad, Npad, N, Mrows;
    const uint16_t* lut16;
    const int* c8_bad;
    const int* tb_eff;         // largest non-rare tB_idx
    const int* tmax;           // [0] largest tA_idx, [1] largest tB_idx
    int64_t esc_limit;         // the byte-cell gate (c8_selected, lut_gemm.cuh): above this many escapes nobody reads D
    int Kpad, t_bias;
    int NT, M, sidx;
    uint32_t nt_magic, zero_word;
    int* D;
};
__global__ void c8_pick_kernel(const unsigned* __restrict__ hist, const int* __restrict__ tmaxB, int64_t elements, int sidx,
                               int* __restrict__ tb_eff)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    int t = *tmaxB;
    if (!sidx) {   // (signed-index layout: the B word hides tB_idx -- no rare class)
        unsigned long long rare = 0;
        while (t > 0 && (rare + hist[t & 63]) * 256ull <= (unsigned long long)elements) { rare += hist[t & 63]; --t; }
    }
    *tb_eff = t;
}
__global__ void __launch_bounds__(256) c8_brare_kernel(const uint32_t* __restrict__ Bw, int64_t Npad, int64_t N, int64_t K, int M,
                                                       const int* __restrict__ tb_eff, unsigned long long* __restrict__ list,
                                                       unsigned cap, unsigned* __restrict__ count)
{
    const int thr = *tb_eff;
    const int64_t total = K * Npad, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x; i0 < total; i0 += stride) {   // warp-uniform trip count
        const int64_t i = i0 + threadIdx.x;
        bool rare = false;
        int64_t k = 0, n = 0;
        if (i < total) {
            if (total < ((int64_t)1 << 32)) { k = (uint32_t)i / (uint32_t)Npad; n = (uint32_t)i - (uint32_t)k * (uint32_t)Npad; }
            else                            { k = i / Npad; n = i - k * Npad; }
            rare = n < N && (int)((__ldg(Bw + i) & 0xFFFFu) >> (M + 1)) > thr;
        }
        const unsigned m = __ballot_sync(0xFFFFFFFFu, rare);
        if (m == 0) continue;
        const unsigned lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(count, __popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (rare) {
            const unsigned pos = base + __popc(m & ((1u << lane) - 1u));
            if (pos < cap) list[pos] = ((unsigned long long)k << 32) | (unsigned long long)(uint32_t)n;
        }
    }
}
__device__ __forceinline__ bool c8_fits(float v) { return v == rintf(v) && fabsf(v) <= 127.0f; }
__global__ void __launch_bounds__(256) c8_correct_kernel(C8CorrectArgs a)
{
    const unsigned nbr = *a.brare_count;
    __shared__ unsigned s_cnt[kEscLists];   // the list counters, fetched together (not one dependent load per list)
    if (threadIdx.x < kEscLists) s_cnt[threadIdx.x] = a.esc_counters[threadIdx.x];
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned tot = 0, ovf = nbr > a.brare_cap;
        for (int l = 0; l < kEscLists; ++l) { tot += s_cnt[l]; ovf |= (s_cnt[l] > a.esc_cap); }
        a.esc_total[0] = tot;
        a.esc_total[1] = ovf;
    }
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int R8 = a.NT - 1 - *a.c8_bad;
    const int thrA = R8 - *a.tb_eff;
    const int tmaxA = a.tmax[0];
    if (tmaxA + a.tmax[1] <= R8) return;   // every pair lies in an all-int8 row: nothing was dropped
    {   // the gate of the byte-cell GEMM, evaluated here too: when it is closed the 16-bit kernel computes and D is never read
        unsigned tot = 0, ovf = nbr > a.brare_cap;
        for (int l = 0; l < kEscLists; ++l) { tot += s_cnt[l]; ovf |= (s_cnt[l] > a.esc_cap); }
        const int dl = tmaxA + a.tmax[1] - a.t_bias + a.M + 1;
        const int lgK = 32 - __clz((a.Kpad > 2 ? a.Kpad : 2) - 1);
        if (ovf || (int64_t)tot > a.esc_limit || dl + lgK > 24) return;
    }
    // Work items are (list entry, 128-wide chunk of the other dimension), dealt round-robin to the warps: thousands of
    // short independent items, each with its four plane loads in flight together (the pass is latency bound).
    constexpr int U = 4;
    // (A escape) x (every column)
    const uint32_t nch = (uint32_t)((a.N + 32 * U - 1) / (32 * U));
    for (int l = 0; l < kEscLists; ++l) {
        const unsigned cnt = min(s_cnt[l], a.esc_cap);
        if (cnt == 0) continue;
        const int64_t items = (int64_t)cnt * nch;
        for (int64_t it = warp; it < items; it += nwarps) {
            const uint32_t e = items < ((int64_t)1 << 32) ? (uint32_t)it / nch : (uint32_t)(it / nch);
            const uint32_t ch = (uint32_t)(it - (int64_t)e * nch);
            const unsigned long long ent = a.esc_list[(size_t)l * a.esc_cap + e];
            const int64_t m = (int64_t)(ent & 0xFFFFFFFFull), k = (int64_t)(ent >> 32);
            const uint32_t aw = __ldg(a.Aw + k * a.Mpad + m);
            uint32_t bw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t n = ((int64_t)ch * U + u) * 32 + lane;
                bw[u] = n < a.N ? __ldg(a.Bw + k * a.Npad + n) : 0u;      // 0: row 0 of the table, all zero cells
            }
            const uint32_t x = aw >> (a.M + 1);
            const int tA = (int)(x - (uint32_t)a.NT * __umulhi(x, a.nt_magic));
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t n = ((int64_t)ch * U + u) * 32 + lane;
                const uint32_t o16 = a.sidx ? bw[u] : (bw[u] & 0xFFFFu);
                // plain layout: the B offset's row is tB_idx; rows up to R8 hold int8 cells only (signed-index layout:
                // the region offset hides tB_idx -- look every pair up)
                if (n >= a.N || (!a.sidx && tA + (int)(o16 >> (a.M + 1)) <= R8)) continue;
                const float v = __uint_as_float((uint32_t)__ldg(a.lut16 + ((aw + o16) >> 1)) << 16);
                if (c8_fits(v)) continue;                                          // the byte table has this cell
                const int sg = a.sidx ? 1 : (((int)bw[u] >> 30) | 1);
                atomicAdd(a.D + m * a.N + n, __float2int_rn(v) * sg);
            }
        }
    }
    // (rare B) x (every row that is not an A escape).  The bulk of the pass: 32-bit index arithmetic (rows < 2^31 is a
    // condition of the variant), eight row chunks per item, the row pointer formed once per item.
    constexpr int U2 = 8;
    const unsigned nb = min(nbr, a.brare_cap);
    const uint32_t mrows = (uint32_t)a.Mrows;
    const uint32_t mch = (mrows + 32 * U2 - 1) / (32 * U2);
    const int64_t items = (int64_t)nb * mch;
    const uint32_t mshift = (uint32_t)(a.M + 1);
    for (int64_t it = warp; it < items; it += nwarps) {
        const uint32_t e = items < ((int64_t)1 << 32) ? (uint32_t)it / mch : (uint32_t)(it / mch);
        const uint32_t m0 = (uint32_t)(it - (int64_t)e * mch) * (32 * U2) + (uint32_t)lane;
        const unsigned long long ent = a.brare_list[e];
        const uint32_t n = (uint32_t)ent, k = (uint32_t)(ent >> 32);
        const uint32_t bw = __ldg(a.Bw + (int64_t)k * a.Npad + n);
        const uint32_t* arow = a.Aw + (int64_t)k * a.Mpad;
        uint32_t aw[U2];
#pragma unroll
        for (int u = 0; u < U2; ++u)    // issued together with the B word: one latency, not two
            aw[u] = (m0 + 32u * u < mrows) ? __ldg(arow + m0 + 32u * u) : a.zero_word;
        const uint32_t o16 = bw & 0xFFFFu;
        const int tB = (int)(o16 >> mshift);
        if (tmaxA + tB <= R8) continue;                      // no row reaches beyond the int8 rows with this element
        const int lo = R8 - tB;                              // rows that matter: lo < tA_idx <= thrA
        const int sg = ((int)bw >> 30) | 1;
        int* dcol = a.D + n;
#pragma unroll
        for (int u = 0; u < U2; ++u) {
            if (aw[u] == a.zero_word) continue;
            const uint32_t x = aw[u] >> mshift;
            const int tA = (int)(x - (uint32_t)a.NT * __umulhi(x, a.nt_magic));
            if (tA > thrA || tA <= lo) continue;             // an escape (done above) / int8 rows only
            const float v = __uint_as_float((uint32_t)__ldg(a.lut16 + ((aw[u] + o16) >> 1)) << 16);
            if (c8_fits(v)) continue;
            atomicAdd(dcol + (int64_t)(m0 + 32u * u) * a.N, __float2int_rn(v) * sg);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Operand predecode: fp32 -> packed table offsets, transposed to [K][rows] so that the GEMM reads
// 128-bit vectors of consecutive rows/columns.  HBM-bound, one pass.
// ------------------------------------------------------------------------------------------------
struct PredecodeArgs {
    const float* src;
    int64_t s_row, s_k;    // element (row, k) at src[row*s_row + k*s_k]   (row = m for A, n for B)
    ASrc asrc;             // A operand only: GEMM rows or implicit im2col
    uint32_t* dst;         // [Kpad][Rpad]
    int64_t rows, K, Rpad, Kpad;
    int M;
    BiasSrc bias;
    int* offgrid;          // raised if any element cannot take the table path
    unsigned long long* zero_count;   // A operand: number of zero elements (sampled; picks the GEMM's k loop), or null
    int* tmax;                        // atomicMax of the operand's table row index (bounds the addends), or null
    // byte-cell variant of the table GEMM (lut_gemm.cuh, A operand only; esc_list == null: off): A elements whose
    // table rows may hold a cell beyond +-127 -- tA_idx > (last all-int8 row) - (largest tB_idx, known because B is
    // decoded first) -- are "escapes": appended as (k << 32 | row) to one of kEscLists lists (block-hashed, so that the
    // list counters are not one contended address); c8_correct_kernel adds the cells the byte table dropped for them
    unsigned long long* esc_list;     // [kEscLists][esc_cap]
    unsigned esc_cap;
    unsigned* esc_counters;           // [kEscLists]
    const int* c8_bad;                // NT - (first row with a cell that does not fit int8), 0 if none
    const int* tmaxB;                 // largest tB_idx among the B elements that are not "rare" (c8_pick_kernel)
    unsigned* hist;                   // B operand: [64] histogram of tB_idx, or null
    __device__ __forceinline__ int thrA(const LutGeom& g) const
    {
        return esc_list ? (g.NT - 1 - *c8_bad) - *tmaxB : 0x7FFFFFFF;
    }
    // warp-aggregated append; every lane of the warp calls it (esc false where there is nothing to report)
    __device__ __forceinline__ void escape(bool esc, int64_t r, int64_t k, unsigned list) const
    {
        const unsigned m = __ballot_sync(0xFFFFFFFFu, esc);
        if (m == 0) return;
        const unsigned lane = threadIdx.x & 31, leader = __ffs(m) - 1;
        unsigned base = 0;
        if (lane == leader) base = atomicAdd(esc_counters + list, __popc(m));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (esc) {
            const unsigned pos = base + __popc(m & ((1u << lane) - 1u));
            if (pos < esc_cap) esc_list[(size_t)list * esc_cap + pos] = ((unsigned long long)k << 32) | (unsigned long long)(uint32_t)r;
        }
    }
};
__device__ __forceinline__ uint32_t predecode_a(const LutGeom& g, float v, int bA, int& off, int& tmax, int thrA, bool& esc)
{
    esc = false;
    const uint32_t bits = __float_as_uint(v);
    const uint32_t mag = bits & 0x7FFFFFFFu;
    if (mag == 0) return 2u * (uint32_t)g.zeroA_off;
    const int M = g.M;
    const int L = (int)(mag >> 23) - 127;
    const uint32_t frac = mag & 0x7FFFFFu;
    const int LA_lo = 1 - bA - M;
    const int tA = L - LA_lo;
    // on grid: <= M fractional bits, at least min_subnorm, not above the top binade, finite/normal fp32
    if ((frac & ((1u << (23 - M)) - 1u)) != 0 || tA < 0 || tA > g.RA || (mag >> 23) == 0 || (mag >> 23) == 255) {
        off = 1;
        return 2u * (uint32_t)g.zeroA_off;
    }
    const uint32_t mA = frac >> (23 - M);
    const uint32_t sa = bits >> 31;
    tmax = max(tmax, tA);
    esc = tA > thrA;
    return 2u * (uint32_t)(g.row_entries * (tA + g.NT * (int)(mA + (sa << M))));
}
__device__ __forceinline__ uint32_t predecode_a(const LutGeom& g, float v, int bA, int& off, int& tmax)
{
    bool esc;
    return predecode_a(g, v, bA, off, tmax, 0x7FFFFFFF, esc);
}

__device__ __forceinline__ uint32_t predecode_b(const LutGeom& g, float v, int bA, int bB, int bR, int& off, int& tmax, int& tbi)
{
    tbi = -1;
    const uint32_t bits = __float_as_uint(v);
    const uint32_t mag = bits & 0x7FFFFFFFu;
    const uint32_t zero_b = g.sidx ? 0u : (kBf16One << 16);   // tB_idx = 0: every product rounds to 0
    if (mag == 0) return zero_b;
    const int M = g.M;
    const int L = (int)(mag >> 23) - 127;
    const uint32_t frac = mag & 0x7FFFFFu;
    int tB = L + bR - bA + 3;                                 // (t + M + 2) - tA_idx
    if ((frac & ((1u << (23 - M)) - 1u)) != 0 || L < 1 - bB - M || (mag >> 23) == 0 || (mag >> 23) == 255 ||
        tB + g.RA > g.NTB - 1 || (1 - bA - M) + (1 - bB - M) < -120) {
        off = 1;
        return zero_b;
    }
    if (tB < -g.RA) tB = -g.RA;
    tmax = max(tmax, tB + g.RA);
    tbi = tB + g.RA;
    const uint32_t mB = frac >> (23 - M);
    const uint32_t boff = 2u * (mB + (uint32_t)(g.row_entries * (tB + g.RA)));
    if (g.sidx) return boff + ((bits >> 31) ? 2u * (uint32_t)g.s_stride : 0u);   // sign of b selects a region
    const uint32_t mult = (bits >> 31) ? kBf16MinusOne : kBf16One;
    return (mult << 16) | boff;
}

// v5 flavour (ofuf_mode 1): operands enter through their CODES (float_to_fpany_absint with clip_OF, v5 :246-281).
// On-grid values only -- exact mantissa, not above the top binade -- anything else raises `off` and the arithmetic
// kernel (which rounds and clips like the reference) recomputes.  A zero is the ordinary code (0, 0) with sign +1
// (sign(v) = v < 0 ? -1 : 1, v5 :106-107): its product is NOT zero in v5.
__device__ __forceinline__ bool v5_code(float v, int M, int E, int bias, int& e, int& m, uint32_t& neg)
{
    const uint32_t bits = __float_as_uint(v);
    const uint32_t mag = bits & 0x7FFFFFFFu;
    e = 0; m = 0; neg = 0;
    if (mag == 0) return true;
    neg = bits >> 31;
    const int ef = (int)(mag >> 23);
    if (ef == 0 || ef == 255) return false;
    const int L = ef - 127, Lmin = 1 - bias;
    const uint32_t frac = mag & 0x7FFFFFu;
    if (L >= Lmin) {
        e = L + bias;
        m = (int)(frac >> (23 - M));
        return (frac & ((1u << (23 - M)) - 1u)) == 0 && e <= (1 << E) - 1;
    }
    const int sh = Lmin - L;                       // sub-normal: |v| = m * 2^(1 - bias - M)
    if (sh > M) return false;
    const uint32_t sig = (1u << 23) | frac;
    m = (int)(sig >> (23 - M + sh));
    return (sig & ((1u << (23 - M + sh)) - 1u)) == 0;
}

__device__ __forceinline__ uint32_t predecode_a_v5(const LutGeom& g, float v, int bA, int& off)
{
    int e, m;
    uint32_t neg;
    if (!v5_code(v, g.M, g.E, bA, e, m, neg)) { off = 1; return 2u * (uint32_t)g.zeroA_off; }
    return 2u * (uint32_t)(g.row_entries * (e + g.NT * (m + (int)(neg << g.M))));
}

__device__ __forceinline__ uint32_t predecode_b_v5(const LutGeom& g, float v, int bA, int bB, int bR, int& off)
{
    int e, m;
    uint32_t neg;
    const uint32_t pad = kBf16One << 16;           // row 0: all zero against every A row
    if (!v5_code(v, g.M, g.E, bB, e, m, neg)) { off = 1; return pad; }
    const int tB = e - (bA + bB - bR);
    if (tB < -g.v5_lo || tB > (1 << g.E) + 3) { off = 1; return pad; }
    const uint32_t boff = 2u * ((uint32_t)m + (uint32_t)(g.row_entries * (tB + g.v5_lo + g.RA + 1)));
    return ((neg ? kBf16MinusOne : kBf16One) << 16) | boff;
}

// 32x32 tile transpose: reads along k (SRC_K_CONTIG) or along rows, writes along rows.
template <bool IS_B, bool SRC_K_CONTIG>
__global__ void __launch_bounds__(256) predecode_kernel(PredecodeArgs a, LutGeom g)
{
    __shared__ uint32_t tile[32][33];
    __shared__ unsigned s_hist[64];   // B operand, byte-cell variant: histogram of tB_idx
    if (IS_B && a.hist) {
        if (threadIdx.x < 64) s_hist[threadIdx.x] = 0;
        __syncthreads();
    }
    const int64_t r0 = (int64_t)blockIdx.x * 32, k0 = (int64_t)blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
    int off = 0, zeros = 0, tmax = 0;
    const int bA = a.bias.A(), bR = a.bias.R();
    const int thrA = IS_B ? 0 : a.thrA(g);
    const unsigned elist = (blockIdx.x * 7u + blockIdx.y * 13u) % kEscLists;
    if (bA <= -64 || bA >= 64 || bR <= -64 || bR >= 64) off = 1;
    if (SRC_K_CONTIG) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t r = r0 + ty + i * 8, k = k0 + tx;
            uint32_t w = IS_B ? (g.sidx ? 0u : (kBf16One << 16)) : 2u * (uint32_t)g.zeroA_off;
            bool esc = false;
            int tbi = -1;
            if (r < a.rows && k < a.K) {
                const float v = IS_B ? a.src[r * a.s_row + k * a.s_k] : a.asrc.load(r, k);
                w = IS_B ? (g.v5 ? predecode_b_v5(g, v, bA, a.bias.B(r), bR, off) : predecode_b(g, v, bA, a.bias.B(r), bR, off, tmax, tbi))
                         : (g.v5 ? predecode_a_v5(g, v, bA, off) : predecode_a(g, v, bA, off, tmax, thrA, esc));
                if (!IS_B) zeros += (v == 0.0f);
            }
            if (!IS_B && a.esc_list) a.escape(esc, r, k, elist);
            if (IS_B && a.hist && tbi >= 0) atomicAdd(&s_hist[tbi & 63], 1u);
            tile[tx][ty + i * 8] = w;   // [k][r]
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t k = k0 + ty + i * 8, r = r0 + tx;
            if (k < a.Kpad && r < a.Rpad) a.dst[k * a.Rpad + r] = tile[ty + i * 8][tx];
        }
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t k = k0 + ty + i * 8, r = r0 + tx;
            uint32_t w = IS_B ? (g.sidx ? 0u : (kBf16One << 16)) : 2u * (uint32_t)g.zeroA_off;
            bool esc = false;
            int tbi = -1;
            if (r < a.rows && k < a.K) {
                const float v = IS_B ? a.src[r * a.s_row + k * a.s_k] : a.asrc.load(r, k);
                w = IS_B ? (g.v5 ? predecode_b_v5(g, v, bA, a.bias.B(r), bR, off) : predecode_b(g, v, bA, a.bias.B(r), bR, off, tmax, tbi))
                         : (g.v5 ? predecode_a_v5(g, v, bA, off) : predecode_a(g, v, bA, off, tmax, thrA, esc));
                if (!IS_B) zeros += (v == 0.0f);
            }
            if (!IS_B && a.esc_list) a.escape(esc, r, k, elist);
            if (IS_B && a.hist && tbi >= 0) atomicAdd(&s_hist[tbi & 63], 1u);
            if (k < a.Kpad && r < a.Rpad) a.dst[k * a.Rpad + r] = w;
        }
    }
    if (off) *a.offgrid = 1;
    if (a.tmax) {
        tmax = __reduce_max_sync(0xFFFFFFFFu, tmax);
        if ((threadIdx.x & 31) == 0 && tmax > *(volatile int*)a.tmax) atomicMax(a.tmax, tmax);
    }
    if (IS_B && a.hist) {
        __syncthreads();
        if (threadIdx.x < 64 && s_hist[threadIdx.x]) atomicAdd(a.hist + threadIdx.x, s_hist[threadIdx.x]);
    }
    // zero statistics of the A operand, sampled: one block in 8 along the rows counts its zeros and its
    // elements (one 64-bit atomic: zeros in the low half, elements in the high half)
    if (!IS_B && a.zero_count && (blockIdx.x & 7) == 0) {
        __shared__ int s_zeros;
        if (threadIdx.x == 0) s_zeros = 0;
        __syncthreads();
        zeros = __reduce_add_sync(0xFFFFFFFFu, zeros);
        if (tx == 0 && zeros) atomicAdd(&s_zeros, zeros);
        __syncthreads();
        if (threadIdx.x == 0) {
            const int64_t nr = min((int64_t)32, a.rows - r0), nk = min((int64_t)32, a.K - k0);
            if (nr > 0 && nk > 0)
                atomicAdd(a.zero_count, (unsigned long long)s_zeros | ((unsigned long long)(nr * nk) << 32));
        }
    }
}

// Conv form (implicit im2col, approx/approx_calculation.py:724-747): one thread owns one output pixel
// (row) and walks 32 consecutive k = (c, ky, kx) with incremental counters -- the row is decomposed
// into (b, ho, wo) ONCE, nothing in the loop divides.  Adjacent threads are adjacent output pixels:
// loads are coalesced for stride-1 convs (and hit L1/L2 for the kh*kw re-reads), stores are 512-byte
// coalesced rows of the k-major plane.  The K-times inflated im2col matrix exists only as this plane.
constexpr int PC_ROWS = 128, PC_KCH = 32;
__global__ void __launch_bounds__(PC_ROWS) predecode_conv_kernel(PredecodeArgs a, LutGeom g)
{
    const ASrc& s = a.asrc;
    const int64_t r = (int64_t)blockIdx.x * PC_ROWS + threadIdx.x;
    const int k0 = blockIdx.y * PC_KCH;
    int off = 0, zeros = 0, tmax = 0;
    const int bA = a.bias.A(), bR = a.bias.R();
    const int thrA = a.thrA(g);
    const unsigned elist = (blockIdx.x * 7u + blockIdx.y * 13u) % kEscLists;
    if (bA <= -64 || bA >= 64 || bR <= -64 || bR >= 64) off = 1;
    const bool valid_r = r < a.rows;
    const int wo = (int)(r % s.Wo);
    const int64_t t2 = r / s.Wo;
    const int ho = (int)(t2 % s.Ho);
    const int64_t b = t2 / s.Ho;
    const int h0 = ho * s.sh - s.ph, w0 = wo * s.sw - s.pw;
    const float* base = s.p + (b * s.C + s.c0) * (int64_t)s.H * s.W;
    int kx = k0 % s.kw, tk = k0 / s.kw;
    int ky = tk % s.kh, c = tk / s.kh;
    const uint32_t zero_word = 2u * (uint32_t)g.zeroA_off;
    uint32_t* dst = a.dst + (int64_t)k0 * a.Rpad + r;
    const int kend = (int)min((int64_t)PC_KCH, a.Kpad - k0);
#pragma unroll 4
    for (int kk = 0; kk < kend; ++kk) {
        uint32_t w = zero_word;
        bool esc = false;
        if (valid_r && k0 + kk < a.K) {
            const int h = h0 + ky * s.dh, wi = w0 + kx * s.dw;
            float v = 0.0f;                                              // F.pad zeros (:727)
            if (h >= 0 && h < s.H && wi >= 0 && wi < s.W) v = s.fq(__ldg(base + ((int64_t)c * s.H + h) * s.W + wi));
            w = g.v5 ? predecode_a_v5(g, v, bA, off) : predecode_a(g, v, bA, off, tmax, thrA, esc);
            zeros += (v == 0.0f);
        }
        if (a.esc_list) a.escape(esc, r, k0 + kk, elist);
        dst[(int64_t)kk * a.Rpad] = w;
        if (++kx == s.kw) { kx = 0; if (++ky == s.kh) { ky = 0; ++c; } }
    }
    if (off) *a.offgrid = 1;
    if (a.tmax) {
        tmax = __reduce_max_sync(0xFFFFFFFFu, tmax);
        if ((threadIdx.x & 31) == 0 && tmax > *(volatile int*)a.tmax) atomicMax(a.tmax, tmax);
    }
    if (a.zero_count && (blockIdx.x & 7) == 0) {   // sampled zero statistics, as in predecode_kernel
        __shared__ int s_zeros;
        if (threadIdx.x == 0) s_zeros = 0;
        __syncthreads();
        zeros = __reduce_add_sync(0xFFFFFFFFu, zeros);
        if ((threadIdx.x & 31) == 0 && zeros) atomicAdd(&s_zeros, zeros);
        __syncthreads();
        if (threadIdx.x == 0) {
            const int64_t nr = min((int64_t)PC_ROWS, a.rows - (int64_t)blockIdx.x * PC_ROWS);
            const int64_t nk = min((int64_t)PC_KCH, a.K - k0);
            if (nr > 0 && nk > 0)
                atomicAdd(a.zero_count, (unsigned long long)s_zeros | ((unsigned long long)(nr * nk) << 32));
        }
    }
}

// Conv form, 1x1 taps with unit stride and no padding (every pointwise conv): im2col is the identity up to the layout
// change NCHW -> [k = channel][row = (b, pixel)], so the pass is elementwise.  One thread converts 4 consecutive pixels
// of one channel (128-bit load, 128-bit store), the fake-quantizer's constants are read once per thread, nothing divides
// per element (predecode_conv_kernel spends ~115 instructions per element on its generic tap arithmetic:
// profiles/r2b_mbv2_predecode_dw_full.txt).  Needs H*W % 4 == 0 and a 16-byte aligned input.
// grid (ceil(Rpad / 4 / 256), Kpad).
__global__ void __launch_bounds__(256) predecode_1x1_kernel(PredecodeArgs a, LutGeom g)
{
    const ASrc& s = a.asrc;
    const int64_t q = (int64_t)blockIdx.x * 256 + threadIdx.x, r = q * 4;
    const int k = blockIdx.y;
    int off = 0, zeros = 0, tmax = 0;
    const int bA = a.bias.A(), bR = a.bias.R();
    const int thrA = a.thrA(g);
    bool e0 = false, e1 = false, e2 = false, e3 = false;
    if (bA <= -64 || bA >= 64 || bR <= -64 || bR >= 64) off = 1;
    const uint32_t zero_word = 2u * (uint32_t)g.zeroA_off;
    const bool live = r < a.rows && k < a.K;
    if (r < a.Rpad) {
        uint4 w = make_uint4(zero_word, zero_word, zero_word, zero_word);
        if (live) {
            const int64_t HW = (int64_t)s.H * s.W;
            int64_t b, p;
            if (a.rows < ((int64_t)1 << 31)) { b = (uint32_t)r / (uint32_t)HW; p = r - b * HW; }
            else                             { b = r / HW; p = r - b * HW; }
            float4 v = __ldg(reinterpret_cast<const float4*>(s.p + ((b * s.C + s.c0 + k) * HW + p)));
            if (s.fq_maxval) {
                const float mx = __ldg(s.fq_maxval), fb = __ldg(s.fq_bias);
                if (fp8_ste_fast_ok(fb)) {
                    const int bi = (int)fb, Mi = (int)s.fq_Mf;
                    v.x = fp8_ste_quant_fast(v.x, mx, bi, Mi, s.fq_sign); v.y = fp8_ste_quant_fast(v.y, mx, bi, Mi, s.fq_sign);
                    v.z = fp8_ste_quant_fast(v.z, mx, bi, Mi, s.fq_sign); v.w = fp8_ste_quant_fast(v.w, mx, bi, Mi, s.fq_sign);
                } else {
                    v.x = fp8_ste_quant(v.x, mx, fb, s.fq_Mf, s.fq_sign); v.y = fp8_ste_quant(v.y, mx, fb, s.fq_Mf, s.fq_sign);
                    v.z = fp8_ste_quant(v.z, mx, fb, s.fq_Mf, s.fq_sign); v.w = fp8_ste_quant(v.w, mx, fb, s.fq_Mf, s.fq_sign);
                }
            }
            if (g.v5) {
                w.x = predecode_a_v5(g, v.x, bA, off); w.y = predecode_a_v5(g, v.y, bA, off);
                w.z = predecode_a_v5(g, v.z, bA, off); w.w = predecode_a_v5(g, v.w, bA, off);
            } else {
                w.x = predecode_a(g, v.x, bA, off, tmax, thrA, e0); w.y = predecode_a(g, v.y, bA, off, tmax, thrA, e1);
                w.z = predecode_a(g, v.z, bA, off, tmax, thrA, e2); w.w = predecode_a(g, v.w, bA, off, tmax, thrA, e3);
            }
            zeros = (v.x == 0.0f) + (v.y == 0.0f) + (v.z == 0.0f) + (v.w == 0.0f);
        }
        *reinterpret_cast<uint4*>(a.dst + (int64_t)k * a.Rpad + r) = w;
    }
    if (off) *a.offgrid = 1;
    if (a.tmax) {
        tmax = __reduce_max_sync(0xFFFFFFFFu, tmax);
        if ((threadIdx.x & 31) == 0 && tmax > *(volatile int*)a.tmax) atomicMax(a.tmax, tmax);
    }
    if (a.esc_list) {
        const unsigned elist = (blockIdx.x * 7u + blockIdx.y * 13u) % kEscLists;
        a.escape(e0, r, k, elist); a.escape(e1, r + 1, k, elist); a.escape(e2, r + 2, k, elist); a.escape(e3, r + 3, k, elist);
    }
    if (a.zero_count && (blockIdx.x & 7) == 0 && (blockIdx.y & 3) == 0) {   // sampled zero statistics
        __shared__ int s_zeros;
        if (threadIdx.x == 0) s_zeros = 0;
        __syncthreads();
        zeros = __reduce_add_sync(0xFFFFFFFFu, zeros);
        if ((threadIdx.x & 31) == 0 && zeros) atomicAdd(&s_zeros, zeros);
        __syncthreads();
        if (threadIdx.x == 0 && k < a.K) {
            const int64_t n = min((int64_t)1024, a.rows - (int64_t)blockIdx.x * 1024);
            if (n > 0) atomicAdd(a.zero_count, (unsigned long long)s_zeros | ((unsigned long long)n << 32));
        }
    }
}

__global__ void __launch_bounds__(256) predecode_plane_kernel(PredecodeArgs a, LutGeom g, uint32_t* __restrict__ Pw,
                                                              int2* __restrict__ kinfo, uint32_t* __restrict__ zero_dst,
                                                              int64_t nB, int64_t Cg)
{
    const ASrc& s = a.asrc;
    const int64_t img = Cg * s.H * s.W, total = nB * img;
    const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int off = 0, zeros = 0, tmax = 0;
    const int bA = a.bias.A(), bR = a.bias.R();
    if (bA <= -64 || bA >= 64 || bR <= -64 || bR >= 64) off = 1;
    if (gid == 0) *zero_dst = 2u * (uint32_t)g.zeroA_off;
    if (gid < a.Kpad) {
        int2 ki = make_int2(0, -30000 & 0xFFFF);   // dy far outside: never in range
        if (gid < a.K) {
            const int kx = (int)(gid % s.kw), t = (int)(gid / s.kw);
            const int ky = t % s.kh, c = t / s.kh;
            ki.x = (c * s.H + ky * s.dh) * s.W + kx * s.dw;
            ki.y = (ky * s.dh) | ((kx * s.dw) << 16);
        }
        kinfo[gid] = ki;
    }
    const int64_t i4 = gid * 4;
    if (i4 < total) {
        // img % 4 == 0 (checked by the host): the four elements are in one image and 16-byte aligned
        const int64_t b = i4 / img, j = i4 - b * img;
        float4 v = __ldg(reinterpret_cast<const float4*>(s.p + (b * s.C + s.c0) * (int64_t)s.H * s.W + j));
        v.x = s.fq(v.x); v.y = s.fq(v.y); v.z = s.fq(v.z); v.w = s.fq(v.w);   // fused input fake-quantizer, if any
        uint4 w;
        w.x = predecode_a(g, v.x, bA, off, tmax);
        w.y = predecode_a(g, v.y, bA, off, tmax);
        w.z = predecode_a(g, v.z, bA, off, tmax);
        w.w = predecode_a(g, v.w, bA, off, tmax);
        zeros = (v.x == 0.0f) + (v.y == 0.0f) + (v.z == 0.0f) + (v.w == 0.0f);
        *reinterpret_cast<uint4*>(Pw + i4) = w;
    }
    if (off) *a.offgrid = 1;
    if (a.tmax) {
        tmax = __reduce_max_sync(0xFFFFFFFFu, tmax);
        if ((threadIdx.x & 31) == 0 && tmax > *(volatile int*)a.tmax) atomicMax(a.tmax, tmax);
    }
    if (a.zero_count && (blockIdx.x & 7) == 0) {   // sampled zero statistics, as in predecode_kernel
        __shared__ int s_zeros;
        if (threadIdx.x == 0) s_zeros = 0;
        __syncthreads();
        zeros = __reduce_add_sync(0xFFFFFFFFu, zeros);
        if ((threadIdx.x & 31) == 0 && zeros) atomicAdd(&s_zeros, zeros);
        __syncthreads();
        if (threadIdx.x == 0) {
            const int64_t n = min((int64_t)1024, total - (int64_t)blockIdx.x * 1024);
            if (n > 0) atomicAdd(a.zero_count, (unsigned long long)s_zeros | ((unsigned long long)n << 32));
        }
    }
}

// stream-K fix-up: one CTA per output tile.  Finds the CTAs whose unit ranges intersect the tile's K range; a
// tile covered by a single CTA was stored by the GEMM itself, otherwise its segments are added in k order.
__global__ void __launch_bounds__(256) streamk_fixup_kernel(const float* __restrict__ P, int G, int64_t tiles, int tiles_n,
                                                            int nk_all, int BM, int BN, float* __restrict__ C, int64_t ldc,
                                                            int64_t Mrows, int64_t N, BiasSrc bias, int M, int qamaa,
                                                            QamaaOp qop, const int* __restrict__ inexact,
                                                            const int* __restrict__ offgrid, Epilogue ep)
{
    if (*inexact != 0 || *offgrid != 0) return;
    const int64_t U = tiles * nk_all, tile = blockIdx.x;
    const int64_t u_lo = tile * nk_all, u_hi = u_lo + nk_all;
    auto start = [&](int64_t g) { return g * U / G; };
    int64_t g0 = u_lo * G / U;
    while (g0 > 0 && start(g0) > u_lo) --g0;
    while (start(g0 + 1) <= u_lo) ++g0;
    if (start(g0 + 1) >= u_hi) return;            // one CTA covered the whole K range: stored directly
    const float unit = pow2f_i(1 - bias.R() - M);
    if (qamaa) qop.begin();
    const int64_t m0 = (tile / tiles_n) * BM, n0 = (tile % tiles_n) * BN;
    // the tile's segments, once per thread (64-bit divisions do not belong in the element loop): CTA g0 reached the
    // tile in a LATER segment of its range (slot 1) unless its range starts exactly here; the CTAs after it start
    // inside the tile (slot 0)
    int64_t g1 = g0 + 1;
    while (g1 < G && start(g1) < u_hi) ++g1;
    const int64_t tile_elems = (int64_t)BM * BN;
    const float* P0 = P + (g0 * 2 + (start(g0) >= u_lo ? 0 : 1)) * tile_elems;
    const int bn_shift = 31 - __clz(BN);          // BN is 64 or 128
    // gridDim.y CTAs share a tile (a lone CTA per tile walked 64 latency-bound iterations per thread)
    for (int e = blockIdx.y * 256 + threadIdx.x; e < BM * BN; e += 256 * gridDim.y) {
        const int r = e >> bn_shift, c = e & (BN - 1);
        if (m0 + r >= Mrows || n0 + c >= N) continue;
        float s = P0[e];
        for (int64_t g = g0 + 1; g < g1; ++g) s += P[g * 2 * tile_elems + e];     // k order
        s *= unit;
        if (qamaa) s = qop.quant(s);
        ep.store(C, ldc, m0 + r, n0 + c, s);
    }
}

// Ordered sum of the split-K slices; does nothing when the exactness certificate failed.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float* __restrict__ P, int splits, float* __restrict__ C,
                                                            int64_t ldc, int64_t Mrows, int64_t N, BiasSrc bias, int M,
                                                            int qamaa, QamaaOp qop, const int* __restrict__ inexact,
                                                            const int* __restrict__ offgrid, Epilogue ep)
{
    if (*inexact != 0 || *offgrid != 0) return;
    const float unit = pow2f_i(1 - bias.R() - M);
    if (qamaa) qop.begin();
    const int64_t total = Mrows * N, stride = (int64_t)gridDim.x * blockDim.x;
    // (row, column) of the element advance incrementally: one 64-bit division per thread, none per element
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t m = i / N, n = i - m * N;
    const int64_t dm = stride / N, dn = stride - dm * N;
    for (; i < total; i += stride) {
        float s = P[i];
        for (int z = 1; z < splits; ++z) s += P[(int64_t)z * total + i];
        s *= unit;
        if (qamaa) s = qop.quant(s);
        ep.store(C, ldc, m, n, s);
        m += dm; n += dn;
        if (n >= N) { n -= N; ++m; }
    }
}

// Same, for a fused epilogue with an NCHW store (epilogue.cuh): 32 rows x 32 columns per CTA, slices read with lanes
// along the columns (contiguous in the partial sums), transposed through shared memory, stored with lanes along the
// rows (= pixels, contiguous in NCHW).  grid (ceil(Mrows / 32), ceil(N / 32)), 256 threads.
__global__ void __launch_bounds__(256) splitk_reduce_nchw_kernel(const float* __restrict__ P, int splits, float* __restrict__ C,
                                                                 int64_t Mrows, int64_t N, BiasSrc bias, int M, int qamaa,
                                                                 QamaaOp qop, const int* __restrict__ inexact,
                                                                 const int* __restrict__ offgrid, Epilogue ep)
{
    if (*inexact != 0 || *offgrid != 0) return;
    __shared__ float tile[32][33];
    const float unit = pow2f_i(1 - bias.R() - M);
    if (qamaa) qop.begin();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int64_t m0 = (int64_t)blockIdx.x * 32, n0 = (int64_t)blockIdx.y * 32, total = Mrows * N;
    const int64_t n = n0 + tx;
    const Epilogue::Col ec = ep.column(n < N ? n : N - 1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = ty + 8 * i;
        const int64_t m = m0 + r;
        float s = 0.0f;
        if (m < Mrows && n < N) {
            const int64_t e = m * N + n;
            s = P[e];
            for (int z = 1; z < splits; ++z) s += P[(int64_t)z * total + e];
            s *= unit;
            if (qamaa) s = qop.quant(s);
            s = ep.apply(s, ec);
        }
        tile[r][tx] = s;
    }
    __syncthreads();
    const int64_t m = m0 + tx;
    if (m < Mrows) {
        const int64_t bimg = m / ep.HW, pix = m - bimg * ep.HW;
        float* obase = C + (bimg * ep.Ctot + ep.c0) * ep.HW + pix;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int c = ty + 8 * i;
            if (n0 + c < N) obase[(n0 + c) * ep.HW] = tile[tx][c];
        }
    }
}

// Tile configurations and the host-side choice between them.
struct LutTileCfg {
    int warps, rpw, cpl, kt, sidx;
    void (*kernel)(LutGemmArgs);
    void (*conv_kernel)(LutGemmArgs);
    void (*sk_kernel)(LutGemmArgs);
    void (*sk_conv_kernel)(LutGemmArgs);
    int BM() const { return warps * rpw; }
    // byte-cell twins (approx_gemm_c8.cu), or null
    void (*kernel8() const)(LutGemmArgs) { void (*k)(LutGemmArgs) = nullptr, (*sk)(LutGemmArgs) = nullptr; lut_c8_kernels(warps, rpw, cpl, kt, sidx, &k, &sk); return k; }
    void (*sk_kernel8() const)(LutGemmArgs) { void (*k)(LutGemmArgs) = nullptr, (*sk)(LutGemmArgs) = nullptr; lut_c8_kernels(warps, rpw, cpl, kt, sidx, &k, &sk); return sk; }
    int BN() const { return 32 * cpl; }
    size_t smem(int lut_bytes) const { return (size_t)((lut_bytes + 127) & ~127) + (size_t)LUT_STAGES * kt * (BM() + BN()) * 4; }
};
#define LUT_TILE(W, R, C, KT, S) \
    {W, R, C, KT, S, approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, false, false>,          \
     approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, true, false>, approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, false, true>, \
     approx_gemm_lut_kernel<W, R, C, KT, (S) != 0, true, true>}
static const LutTileCfg kLutTiles[] = {
    LUT_TILE(8, 16, 4, 16, 0),    // 128 x 128
    LUT_TILE(8, 32, 2, 16, 0),    // 256 x  64   (N <= 64)
    LUT_TILE(4, 16, 4, 16, 0),    //  64 x 128   (few rows)
    LUT_TILE(4, 32, 2, 16, 0),    // 128 x  64
    LUT_TILE(4, 16, 2, 16, 0),    //  64 x  64
    LUT_TILE(16, 16, 4, 16, 0),   // 256 x 128   (large tables: one CTA per SM)
    // signed-index tables (three sign regions, up to 176 KB for E2M5)
    LUT_TILE(16, 16, 4, 8, 1),    // 256 x 128, shallow k tile
    LUT_TILE(8, 16, 4, 8, 1),     // 128 x 128
    LUT_TILE(8, 16, 4, 16, 1),
    LUT_TILE(4, 16, 4, 16, 1),
    LUT_TILE(4, 16, 2, 16, 1),
};

// Largest split-K factor the workspace reserves partial-sum room for: up to 8 slices, 256 MB in all (HBM is
// 180 GB; the partial sums cost one extra write + read of S*Mrows*N*4 bytes, a few % of the GEMM's time).
static int max_splits(int64_t Mrows, int64_t N)
{
    const int64_t el = Mrows * N;
    if (el <= 0 || el > ((int64_t)1 << 24)) return 1;
    int64_t cap = ((int64_t)64 << 20) / el;
    if (const char* e = getenv("B200_LUT_PARTIAL_MB")) cap = ((int64_t)atoi(e) << 18) / el;   // experiments
    return (int)std::max<int64_t>(1, std::min<int64_t>(8, cap));
}
static size_t conv_plane_bytes(int64_t plane_words, int64_t Kpad)
{
    return (size_t)round_up(plane_words, 4) * 4 + (size_t)Kpad * 8 + 256;
}
// stream-K partial tiles: 2 slots per persistent CTA, at most 2 * 592 * 8192 (or 2 * 296 * 16384) floats
constexpr size_t kStreamKBytes = (size_t)40 << 20;
static size_t splitk_partial_bytes(int64_t Mrows, int64_t N)
{
    const int64_t el = Mrows * N;
    if (el <= 0 || el > ((int64_t)1 << 24)) return 0;
    const int s = max_splits(Mrows, N);
    return std::max<size_t>(kStreamKBytes, s > 1 ? round_up((size_t)s * Mrows * N * 4, 256) : 0);
}
struct LutTileCfg;
static int resident_ctas(const LutTileCfg& c, int lut_bytes);

static int resident_ctas(const LutTileCfg& c, int lut_bytes)
{
    const size_t sm = c.smem(lut_bytes) + 1024;
    if (sm > 227 * 1024) return 0;
    int r = (int)((227 * 1024) / sm);
    const int acc_regs = c.rpw * c.cpl >= 64 ? 128 : 80;
    r = std::min(r, 65536 / (acc_regs * 32 * c.warps));
    r = std::min(r, 64 / c.warps);
    return r;
}

// Cost model (in table lookups per SM): an SM's lookup rate saturates with 8 resident warps; a CTA does
// BM*BN lookups per k; CTAs are dealt round-robin over the SMs, `r` resident at a time.  Splitting k by S
// multiplies the CTA count and divides each CTA's k range; it costs a pipeline prologue/epilogue per CTA
// and two more launches (the ordered reduction and the gated unsplit fallback).
struct LutCand {
    double cost;
    const LutTileCfg* cfg;
    int S;
};
static const LutTileCfg* choose_lut_tile(int64_t Mrows, int64_t N, int64_t Kpad, int lut_bytes, int sidx, int sms,
                                         int max_s, int* splits, std::vector<LutCand>* all = nullptr, bool* forced = nullptr)
{
    if (forced) *forced = getenv("B200_LUT_SPLITK") || getenv("B200_LUT_TILE");
    const LutTileCfg* best = nullptr;
    double best_cost = 0;
    *splits = 1;
    int forced_s = 0;
    double rate8 = 0.7;
    if (const char* e = getenv("B200_LUT_RATE8")) rate8 = atof(e);   // experiments
    if (const char* e = getenv("B200_LUT_SPLITK")) forced_s = std::max(1, std::min(atoi(e), max_s));   // experiments / tests
    if (const char* e = getenv("B200_LUT_TILE")) {   // experiments: force a tile configuration by index
        const int i = atoi(e);
        if (i >= 0 && i < (int)(sizeof(kLutTiles) / sizeof(kLutTiles[0])) && kLutTiles[i].sidx == sidx &&
            kLutTiles[i].smem(lut_bytes) + 1024 <= 227 * 1024) {
            if (forced_s && Kpad / kLutTiles[i].kt >= forced_s) *splits = forced_s;
            return &kLutTiles[i];
        }
    }
    for (int pass = 0; pass < 2 && !best; ++pass, forced_s = 0)
    for (const LutTileCfg& c : kLutTiles) {
        if (c.sidx != sidx) continue;
        const int r = resident_ctas(c, lut_bytes);
        if (r < 1) continue;
        const int64_t nk = Kpad / c.kt;
        const double T = (double)c.BM() * c.BN() * c.kt * (c.cpl == 2 ? 1.06 : 1.0) * (c.kt < 16 ? 1.05 : 1.0);
        // lookup rate of an SM against resident warps (latency hiding): measured ~0.7 of the peak rate with 8
        // warps (one 128x128 CTA per SM), ~1.0 from 16 warps on
        auto rate = [&](int ctas) {
            const double w = ctas * c.warps;
            return w <= 8.0 ? rate8 * w / 8.0 : std::min(1.0, rate8 + (1.0 - rate8) * (w - 8.0) / 8.0);
        };
        for (int S = 1; S <= max_s; ++S) {
            if (forced_s && S != forced_s) continue;
            if (S > 1 && nk / S < 4) break;
            // two full waves of whole tiles already balance well: measured on 4096^3, splitting then only adds
            // the partial-sum traffic (6.36 vs 6.43 T MAC/s)
            if (S > 1 && !forced_s && ceil_div(Mrows, c.BM()) * ceil_div(N, c.BN()) >= (int64_t)2 * sms * r) break;
            const int64_t nc = ceil_div(Mrows, c.BM()) * ceil_div(N, c.BN()) * S;
            const int64_t full = nc / ((int64_t)sms * r), rem = nc - full * sms * r;
            const int rem_per_sm = (int)ceil_div(rem, sms);
            const double Tc = T * ((double)ceil_div(nk, S) + 2.0);
            double cost = full * r * Tc / rate(r);
            if (rem_per_sm > 0) cost += rem_per_sm * Tc / rate(rem_per_sm);
            if (S > 1) cost += 3.0e5 + 4.0 * (double)S * Mrows * N / sms;
            if (!best || cost < best_cost) { best = &c; best_cost = cost; *splits = S; }
            if (all) all->push_back(LutCand{cost, &c, S});
        }
        // stream-K (S == 0): the (tile, k-tile) units dealt evenly over sms * r persistent CTAs -- no tile
        // quantisation; costs the partial-tile traffic and the fix-up launch.  Only a measured candidate.
        if (all && max_s > 1 && !forced_s) {
            const int64_t tiles = ceil_div(Mrows, c.BM()) * ceil_div(N, c.BN()), G = (int64_t)sms * r;
            if (tiles * nk >= 4 * G && tiles < 8 * G && (size_t)(2 * G) * c.BM() * c.BN() * 4 <= kStreamKBytes) {
                const double cost = (double)tiles * T * ((double)nk + 2.0) / sms / rate(r) + 2.0 * T + 3.0e5 +
                                    8.0 * (double)G * c.BM() * c.BN() / sms;
                all->push_back(LutCand{cost, &c, 0});
            }
        }
    }
    return best;
}

// ------------------------------------------------------------------------------------------------
// Arithmetic ("generic") kernel: pair_generic() per operand pair on the pair_gemm skeleton.
// ------------------------------------------------------------------------------------------------
struct ApproxPairOp {
    typedef FmtParam Col;
    BiasSrc bias;
    PairCfg cfg;
    const signed char* T;   // device table (4^M int8) or nullptr (= zeros)
    FmtParam pA, pR;
    __device__ __forceinline__ void begin()
    {
        pA = make_param(cfg.E, cfg.M, bias.A(), cfg.quirk);
        pR = make_param(cfg.E, cfg.M, bias.R(), cfg.quirk);
    }
    __device__ __forceinline__ Col column(int64_t n) const { return make_param(cfg.E, cfg.M, bias.B(n), cfg.quirk); }
    __device__ __forceinline__ float pair(float a, float b, const Col& pB) const
    {
        return T ? pair_any(cfg, pA, pB, pR, T, a, b) : pair_any<signed char>(cfg, pA, pB, pR, nullptr, a, b);
    }
    __device__ __forceinline__ float finish(float acc, const Col&) const { return acc; }
};

// ------------------------------------------------------------------------------------------------
// Host side: table cache, dispatch, C ABI.
// ------------------------------------------------------------------------------------------------
struct LutEntry {
    LutGeom geom;
    uint16_t* d_lut = nullptr;
    int lut_bytes = 0;
    LutFlags flags{0, 0};
    int8_t* d_lut8 = nullptr;     // byte-cell companion (M >= 4, approx_v9 flavour), with d_bad[0] = NT - first unfit row
    int lut8_bytes = 0;
    int* d_bad = nullptr;
};

static std::mutex g_cache_mu;
static std::map<std::tuple<int, const void*>, signed char*> g_dev_tables;            // (device, host table) -> device copy
static std::map<std::tuple<int, int, int, const void*, int, int>, LutEntry> g_luts;   // (device,E,M,table,with_approx,clip)

int device_table(const signed char* host, int M, const signed char** dev_out)
{
    *dev_out = nullptr;
    if (!host) return B200_OK;
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_cache_mu);
    auto key = std::make_tuple(dev, (const void*)host);
    auto it = g_dev_tables.find(key);
    if (it == g_dev_tables.end()) {
        signed char* d = nullptr;
        const size_t n = (size_t)(1 << M) * (1 << M);
        B200_CUDA_TRY(cudaMalloc(&d, n));
        B200_CUDA_TRY(cudaMemcpy(d, host, n, cudaMemcpyHostToDevice));
        it = g_dev_tables.emplace(key, d).first;
    }
    *dev_out = it->second;
    return B200_OK;
}

// Formats whose table GEMM has a byte-cell variant: the wide mantissas, whose 16-bit cells spread a warp's lookups over
// more words than there are banks.  B200_LUT_C8=0 switches it off (before / after measurements).
static bool c8_format(int M)
{
    static const bool on = !(getenv("B200_LUT_C8") && atoi(getenv("B200_LUT_C8")) == 0);
    return on && M >= 4;
}

// Builds (once per configuration and device) the pair table of lut.cuh.  First use synchronises.
struct V5Flags {
    int on, sim, of_opt, uf_opt;
};

static int build_lut(int E, int M, const signed char* dev_table, int with_approx, int clip_OF, bool sidx, bool quirk,
                     LutEntry* e, const V5Flags* v5 = nullptr)
{
    e->geom = (v5 && v5->on) ? make_lut_geom_v5(E, M) : make_lut_geom(E, M, sidx);
    e->lut_bytes = (int)round_up((int64_t)e->geom.total_entries * 2, 16);
    LutFlags* d_flags = nullptr;
    B200_CUDA_TRY(cudaMalloc(&e->d_lut, e->lut_bytes));
    B200_CUDA_TRY(cudaMalloc(&d_flags, sizeof(LutFlags)));
    B200_CUDA_TRY(cudaMemset(e->d_lut, 0, e->lut_bytes));
    B200_CUDA_TRY(cudaMemset(d_flags, 0, sizeof(LutFlags)));
    if (v5 && v5->on)
        lut_build_v5_kernel<<<(int)ceil_div(e->geom.total_entries, 256), 256>>>(e->geom, dev_table, v5->sim, v5->of_opt,
                                                                               v5->uf_opt, e->d_lut, d_flags);
    else
        lut_build_kernel<<<(int)ceil_div(e->geom.total_entries, 256), 256>>>(e->geom, dev_table, dev_table != nullptr,
                                                                            with_approx, clip_OF, quirk ? 1 : 0, e->d_lut,
                                                                            d_flags);
    B200_LAUNCH_CHECK("lut_build_kernel");
    B200_CUDA_TRY(cudaMemcpy(&e->flags, d_flags, sizeof(LutFlags), cudaMemcpyDeviceToHost));
    cudaFree(d_flags);
    if (c8_format(M) && !(v5 && v5->on) && !quirk) {
        e->lut8_bytes = (int)round_up((int64_t)e->geom.total_entries, 16);
        B200_CUDA_TRY(cudaMalloc(&e->d_lut8, e->lut8_bytes));
        B200_CUDA_TRY(cudaMalloc(&e->d_bad, sizeof(int)));
        B200_CUDA_TRY(cudaMemset(e->d_bad, 0, sizeof(int)));
        lut8_build_kernel<<<(int)ceil_div(e->lut8_bytes, 256), 256>>>(e->geom, e->d_lut, e->d_lut8, e->lut8_bytes, e->d_bad);
        B200_LAUNCH_CHECK("lut8_build_kernel");
        B200_CUDA_TRY(cudaDeviceSynchronize());
    }
    return B200_OK;
}

static int get_lut(int E, int M, const signed char* host_table, const signed char* dev_table, int with_approx,
                   int clip_OF, const LutEntry** out, bool quirk = false, const V5Flags* v5 = nullptr)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_cache_mu);
    const bool is_v5 = v5 && v5->on;
    const int mode = is_v5 ? (4 + 8 * v5->sim + 16 * v5->of_opt + 32 * v5->uf_opt) : (clip_OF + (quirk ? 2 : 0));
    auto key = std::make_tuple(dev, E, M, (const void*)host_table, is_v5 ? 0 : with_approx, mode);
    auto it = g_luts.find(key);
    if (it == g_luts.end()) {
        LutEntry e;
        if (int rc = build_lut(E, M, dev_table, with_approx, clip_OF, quirk, quirk, &e, v5)) return rc;
        if (!is_v5 && !quirk && e.flags.sign_quirk && !e.flags.unrepresentable) {
            // not odd in the operand signs: rebuild with both signs in the index (lut.cuh)
            cudaFree(e.d_lut);
            if (e.d_lut8) cudaFree(e.d_lut8);
            if (e.d_bad) cudaFree(e.d_bad);
            e = LutEntry();
            if (int rc = build_lut(E, M, dev_table, with_approx, clip_OF, true, false, &e)) return rc;
        }
        it = g_luts.emplace(key, e).first;
    }
    *out = &it->second;
    return B200_OK;
}


// The single-column table (signed-index layout, quirk numerics) for conv.cu's depthwise kernel.
int get_quirk_lut(int E, int M, const signed char* host_table, const signed char* dev_table, int with_approx,
                  LutGeom* geom, const uint16_t** d_lut, int* lut_bytes, int* usable)
{
    const LutEntry* e = nullptr;
    if (int rc = get_lut(E, M, host_table, dev_table, with_approx, 0, &e, true)) return rc;
    *geom = e->geom;
    *d_lut = e->d_lut;
    *lut_bytes = e->lut_bytes;
    *usable = e->flags.unrepresentable == 0;
    return B200_OK;
}

// 2-D tensor map over a k-major uint32 plane [outer = Kpad][inner = Rpad], box {box_inner, box_outer}, no swizzle
// (the kernel reads the tile as plain rows).  cuTensorMapEncodeTiled through the runtime's driver entry point: the
// library does not link libcuda.
typedef CUresult (*LutEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static int lut_plane_map(CUtensorMap* map, const uint32_t* base, int64_t inner, int64_t outer, int box_inner, int box_outer)
{
    static std::mutex mu;
    static LutEncodeTiledFn fn = nullptr;
    {
        std::lock_guard<std::mutex> lk(mu);
        if (!fn) {
            void* p = nullptr;
            cudaDriverEntryPointQueryResult q;
            B200_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
            if (!p || q != cudaDriverEntryPointSuccess) { set_error("cuTensorMapEncodeTiled is not available in this driver"); return B200_ERR_CUDA; }
            fn = reinterpret_cast<LutEncodeTiledFn>(p);
        }
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
    const cuuint64_t gstride[1] = {(cuuint64_t)inner * 4};
    const cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint32_t*>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (code " + std::to_string((int)r) + ")"); return B200_ERR_CUDA; }
    return B200_OK;
}

// Optional per-launch timing of the dominant kernel (bench.py's roofline leg): CUDA events recorded
// on the launching stream right around the table GEMM.
static std::mutex g_prof_mu;
static bool g_prof_on = false;
static std::vector<std::pair<cudaEvent_t, cudaEvent_t>> g_prof_events;

static void prof_record(bool start, cudaStream_t st)
{
    std::lock_guard<std::mutex> lk(g_prof_mu);
    if (!g_prof_on) return;
    if (start) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        g_prof_events.emplace_back(e0, e1);
        cudaEventRecord(e0, st);
    } else if (!g_prof_events.empty()) {
        cudaEventRecord(g_prof_events.back().second, st);
    }
}

// Measured choice of (tile, split-K factor).  The cost model above ranks the candidates; the first call for a
// new problem shape times the best few on the caller's own operands (CUDA events on the caller's stream, one
// host synchronisation per candidate) and caches the winner.  Every candidate computes the same bits (split-K
// runs under the exactness certificate), so the choice never changes a result.  Skipped while the stream is
// being captured into a CUDA graph, for small problems, and when B200_LUT_TILE / B200_LUT_SPLITK force a
// configuration.  OFF by default (no entry point synchronises with the host); b200_set_autotune(1) or
// B200_AUTOTUNE=1 opts in.
struct TuneKey {
    int64_t Mrows, N, Kpad;
    int lut_bytes, sidx, gather, max_s, qamaa;
    bool operator<(const TuneKey& o) const
    {
        return std::tie(Mrows, N, Kpad, lut_bytes, sidx, gather, max_s, qamaa) <
               std::tie(o.Mrows, o.N, o.Kpad, o.lut_bytes, o.sidx, o.gather, o.max_s, o.qamaa);
    }
};
static std::map<TuneKey, std::pair<const LutTileCfg*, int>> g_tuned;
static std::mutex g_tuned_mu;
static std::atomic<int> g_autotune{-1};
static std::atomic<uint64_t> g_tune_runs{0};
static bool autotune_on()
{
    int v = g_autotune.load(std::memory_order_relaxed);
    if (v < 0) {
        const char* e = getenv("B200_AUTOTUNE");
        v = e ? (atoi(e) != 0) : 0;   // off unless asked for: "no host sync inside" is the default contract
        g_autotune.store(v, std::memory_order_relaxed);
    }
    return v != 0;
}
constexpr int kTuneCandidates = 6;
constexpr int64_t kTuneMinMacs = (int64_t)1 << 26;

constexpr int64_t kPadM = 256, kPadN = 128, kPadK = 32;
constexpr size_t kFlagBytes = 256;   // ints: [0] off-grid, [2..3] zero statistics, [4..5] row maxima, [6] split-K inexact,
                                     // [8] escapes, [9] escape-list overflow, [10] qamaa: byte-cell table's "bad row" word,
                                     // [11] byte cells ran, [12] largest non-rare tB_idx, [13] rare B elements, [16] qamaa: result bias, [32..63] escape-list counters
constexpr size_t kFlagClearBytes = kFlagBytes;
// byte-cell variant, extra workspace: [histogram of tB_idx, 256 B][correction sums D[Mrows][N]][rare-B list]
// [escape lists of A: room for 1/8 of the A elements, 8 bytes each]
static unsigned esc_list_cap(int64_t Mpad, int64_t Kpad) { return (unsigned)std::max<int64_t>(64, Mpad * Kpad / 8 / kEscLists); }
static size_t esc_list_bytes(int64_t Mpad, int64_t Kpad) { return (size_t)esc_list_cap(Mpad, Kpad) * kEscLists * 8; }
static unsigned brare_cap(int64_t Npad, int64_t Kpad) { return (unsigned)std::max<int64_t>(64, Npad * Kpad / 128); }
static size_t c8_D_bytes(int64_t Mrows, int64_t N) { return (size_t)round_up(Mrows * N * 4, 256); }
static size_t c8_extra_bytes(int64_t Mrows, int64_t N, int64_t Mpad, int64_t Npad, int64_t Kpad)
{
    return 256 + c8_D_bytes(Mrows, N) + (size_t)brare_cap(Npad, Kpad) * 8 + esc_list_bytes(Mpad, Kpad);
}
// Prepared right operand (b200_approx_params.prepared_b): [256 B: image of the flag words the B passes produce]
// [decoded plane Bw[Kpad][Npad]][rare-B list].  fill: the B passes write into it; ready: they are skipped and the
// flag words restored.
struct PreparedB {
    uint8_t* base = nullptr;
    bool ready = false;
};
static size_t prepared_b_bytes(int64_t K, int64_t N)
{
    if (K <= 0 || N <= 0) return 0;
    const int64_t Npad = round_up(N, kPadN), Kpad = round_up(K, kPadK);
    return 256 + (size_t)Kpad * Npad * 4 + (size_t)brare_cap(Npad, Kpad) * 8;
}
// flag words of the B passes: [0] off-grid (OR-ed), [5] largest tB_idx, [12] largest non-rare tB_idx, [13] rare count
// restore (64 threads) also CLEARS the rest of the flag block: the call that reuses a prepared B needs no memset node
__global__ void b_meta_kernel(int* __restrict__ flag, int* __restrict__ meta, int save)
{
    if (blockIdx.x != 0) return;
    if (save) {
        if (threadIdx.x == 0) { meta[0] = flag[0]; meta[1] = flag[5]; meta[2] = flag[12]; meta[3] = flag[13]; }
        return;
    }
    const int t = threadIdx.x;
    if (t >= 64) return;            // kFlagBytes / 4 words
    int v = 0;
    if (t == 0) v = meta[0] != 0;
    else if (t == 5) v = meta[1];
    else if (t == 12) v = meta[2];
    else if (t == 13) v = meta[3];
    flag[t] = v;
}

struct C8Src {
    const int8_t* lut8 = nullptr;
    int lut8_bytes = 0;
    const int* bad = nullptr;
    uint8_t* extra = nullptr;   // c8_extra_bytes
};
static thread_local int g_last_splits = 1;

// Predecode of both operands + the table GEMM launch, shared by the approx path and the
// quantize_after_mult_and_add path (whose table and result bias live in the workspace).
static int run_lut_gemm(const ASrc& A, const float* B, int64_t sb_k, int64_t sb_n, float* C, int64_t ldc,
                        int64_t Mrows, int64_t K, int64_t N, const BiasSrc& bias, const LutGeom& geom,
                        const uint16_t* d_lut, int lut_bytes, int* flag, uint32_t* Aw, uint32_t* Bw, int64_t Mpad,
                        int64_t Npad, int64_t Kpad, const QamaaOp* qop, float* partial, void* conv_ws, size_t conv_ws_bytes,
                        cudaStream_t st, const Epilogue& ep = Epilogue{}, const C8Src& c8src = C8Src{},
                        const PreparedB& prep = PreparedB{})
{
    if (prep.base) Bw = reinterpret_cast<uint32_t*>(prep.base + 256);   // the decoded right operand lives with the caller
    const int M = geom.M;
    static PerDeviceOnce once;
    int sms = 0;
    if (int rc = once.get(&sms, []() -> int {
            for (const LutTileCfg& c : kLutTiles) {
                B200_CUDA_TRY(cudaFuncSetAttribute(c.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                B200_CUDA_TRY(cudaFuncSetAttribute(c.conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                B200_CUDA_TRY(cudaFuncSetAttribute(c.sk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
                B200_CUDA_TRY(cudaFuncSetAttribute(c.sk_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
            }
            return lut_c8_set_attributes();
        }))
        return rc;
    int S = 1;
    const int max_s = partial ? max_splits(Mrows, N) : 1;
    std::vector<LutCand> cands;
    bool forced = false;
    const LutTileCfg* tc = choose_lut_tile(Mrows, N, Kpad, lut_bytes, geom.sidx, sms, max_s, &S, &cands, &forced);
    if (!tc) { set_error("b200_approx_gemm: pair table does not fit shared memory"); return B200_ERR_ARGUMENT; }
    // byte-cell variant: needs the companion table, the flag map and pre-gathered planes (decided below for convs)
    const bool c8_inputs = c8src.lut8 && c8src.extra && c8src.bad && !geom.v5 && Mpad < ((int64_t)1 << 31) &&
                           Mpad * Kpad / 8 < ((int64_t)1 << 31);
    PredecodeArgs pa;
    pa.tmax = (partial || c8_inputs || prep.base) ? flag + 4 : nullptr;
    pa.src = nullptr; pa.s_row = 0; pa.s_k = 1; pa.asrc = A; pa.dst = Aw; pa.rows = Mrows; pa.K = K; pa.Rpad = Mpad;
    pa.Kpad = Kpad; pa.M = M; pa.bias = bias; pa.offgrid = flag;
    pa.esc_list = nullptr; pa.esc_cap = esc_list_cap(Mpad, Kpad); pa.esc_counters = reinterpret_cast<unsigned*>(flag + 32);
    pa.c8_bad = c8src.bad; pa.tmaxB = flag + 12; pa.hist = nullptr;
    unsigned* c8_hist = reinterpret_cast<unsigned*>(c8src.extra);
    int* c8_D = c8_inputs ? reinterpret_cast<int*>(c8src.extra + 256) : nullptr;
    unsigned long long* brare_list =
        prep.base ? reinterpret_cast<unsigned long long*>(prep.base + 256 + (size_t)Kpad * Npad * 4)
                  : (c8_inputs ? reinterpret_cast<unsigned long long*>(c8src.extra + 256 + c8_D_bytes(Mrows, N)) : nullptr);
    if (c8_inputs)   // histogram + correction sums start at zero
        B200_CUDA_TRY(cudaMemsetAsync(c8src.extra, 0, 256 + (size_t)Mrows * N * 4, st));
    // v5 flavour: a zero operand is an ordinary code (its product is not zero), so there is nothing to skip
    pa.zero_count = geom.v5 ? nullptr : reinterpret_cast<unsigned long long*>(flag + 2);
    // B first: the A pass of the byte-cell variant compares against the largest tB_idx
    if (prep.base && prep.ready) {
        b_meta_kernel<<<1, 64, 0, st>>>(flag, reinterpret_cast<int*>(prep.base), 0);   // clears the flag block too
        B200_LAUNCH_CHECK("b_meta_kernel(restore)");
    } else {
        PredecodeArgs pb = pa;
        pb.zero_count = nullptr;
        pb.tmax = pa.tmax ? flag + 5 : nullptr;
        pb.src = B; pb.s_row = sb_n; pb.s_k = sb_k; pb.dst = Bw; pb.rows = N; pb.Rpad = Npad;
        pb.hist = (c8_inputs && !geom.sidx) ? c8_hist : nullptr;
        const dim3 pgrid((unsigned)(Npad / 32), (unsigned)(Kpad / 32));
        if (sb_k == 1 && sb_n != 1) predecode_kernel<true, true><<<pgrid, 256, 0, st>>>(pb, geom);
        else                        predecode_kernel<true, false><<<pgrid, 256, 0, st>>>(pb, geom);
        B200_LAUNCH_CHECK("predecode_kernel<B>");
        if (c8_inputs) {
            c8_pick_kernel<<<1, 32, 0, st>>>(c8_hist, flag + 5, K * N, geom.sidx, flag + 12);
            B200_LAUNCH_CHECK("c8_pick_kernel");
            if (!geom.sidx) {
                c8_brare_kernel<<<(unsigned)std::min<int64_t>(ceil_div(K * Npad, 256), 16 * sms), 256, 0, st>>>(
                    Bw, Npad, N, K, M, flag + 12, brare_list, brare_cap(Npad, Kpad), reinterpret_cast<unsigned*>(flag + 13));
                B200_LAUNCH_CHECK("c8_brare_kernel");
            }
        }
        if (prep.base) {
            b_meta_kernel<<<1, 32, 0, st>>>(flag, reinterpret_cast<int*>(prep.base), 1);
            B200_LAUNCH_CHECK("b_meta_kernel(save)");
        }
    }
    // Conv form: decode the input once and let the GEMM gather (needs 16-byte-aligned images and the extra
    // workspace b200_approx_conv2d_workspace_bytes reserves); else write the im2col plane from the NCHW tensor
    // (contiguous along the output pixel, i.e. along rows: read and write row-fastest).
    // GEMM form: rows are k-contiguous (transpose through shared memory).
    ConvPlane cp{};
    bool gather = false;
    if (A.conv) {
        const int64_t Cg = K / ((int64_t)A.kh * A.kw), img = Cg * A.H * A.W;
        const int64_t nB = Mrows / ((int64_t)A.Ho * A.Wo), total = nB * img;
        const size_t need = conv_plane_bytes(total, Kpad);
        // The gather costs the GEMM ~4-12 % (4-byte copies, address arithmetic) and saves the kh*kw-fold plane
        // write of the predecode pass, which is amortised over the N columns: measured on ResNet-18 it wins
        // for 3x3 / 7x7 taps up to N ~ 256 and loses for 1x1 taps (nothing to save) and wide layers.
        const char* e = getenv("B200_CONV_GATHER");   // experiments / tests: 0 = never, 1 = heuristic, 2 = whenever possible
        const int mode = e ? atoi(e) : 1;
        const bool pays = mode >= 2 || (A.kh * A.kw >= 4 && N <= 256);
        // (v5 flavour: padding taps are REAL zeros of the reference's im2col matrix and must decode as the code
        // (0, 0), which the gather's shared "zero word" cannot tell from k / row padding: im2col plane only)
        gather = !geom.v5 && mode != 0 && pays && conv_ws && conv_ws_bytes >= need && img % 4 == 0 && total < ((int64_t)1 << 31) &&
                 ((uintptr_t)A.p & 15) == 0 && A.H < 16384 && A.W < 16384 && nB * A.Ho * A.Wo == Mrows;
    }
    const bool c8 = c8_inputs && !gather;
    if (c8)   // the escape lists follow the workspace's own rare-B list room (unused when the caller keeps B prepared)
        pa.esc_list = reinterpret_cast<unsigned long long*>(c8src.extra + 256 + c8_D_bytes(Mrows, N)) + brare_cap(Npad, Kpad);
    if (A.conv) {
        const int64_t Cg = K / ((int64_t)A.kh * A.kw), img = Cg * A.H * A.W;
        const int64_t nB = Mrows / ((int64_t)A.Ho * A.Wo), total = nB * img;
        if (gather) {
            uint32_t* Pw = reinterpret_cast<uint32_t*>(conv_ws);
            int2* kinfo = reinterpret_cast<int2*>(Pw + round_up(total, 4));
            uint32_t* zsrc = reinterpret_cast<uint32_t*>(kinfo + Kpad);
            cp.Pw = Pw; cp.kinfo = kinfo; cp.zero_src = zsrc; cp.img_words = img;
            cp.H = A.H; cp.W = A.W; cp.Ho = A.Ho; cp.Wo = A.Wo; cp.sh = A.sh; cp.sw = A.sw; cp.ph = A.ph; cp.pw = A.pw;
            const int64_t threads = std::max<int64_t>(ceil_div(total, 4), Kpad);
            predecode_plane_kernel<<<(unsigned)ceil_div(threads, 256), 256, 0, st>>>(pa, geom, Pw, kinfo, zsrc, nB, Cg);
        } else if (A.kh == 1 && A.kw == 1 && A.sh == 1 && A.sw == 1 && A.ph == 0 && A.pw == 0 && ((int64_t)A.H * A.W) % 4 == 0 &&
                   ((uintptr_t)A.p & 15) == 0 && A.Ho == A.H && A.Wo == A.W && Kpad <= 65535 &&
                   !(getenv("B200_PREDECODE_1X1") && atoi(getenv("B200_PREDECODE_1X1")) == 0)) {
            predecode_1x1_kernel<<<dim3((unsigned)ceil_div(Mpad / 4, 256), (unsigned)Kpad), 256, 0, st>>>(pa, geom);
        } else {
            predecode_conv_kernel<<<dim3((unsigned)(Mpad / PC_ROWS), (unsigned)ceil_div(Kpad, PC_KCH)), PC_ROWS, 0, st>>>(pa, geom);
        }
    } else {
        predecode_kernel<false, true><<<dim3((unsigned)(Mpad / 32), (unsigned)(Kpad / 32)), 256, 0, st>>>(pa, geom);
    }
    B200_LAUNCH_CHECK("predecode_kernel<A>");
    const uint32_t nt_magic = (uint32_t)(((uint64_t)1 << 32) / (uint64_t)geom.NT) + 1u;
    if (c8) {
        C8CorrectArgs ca;
        ca.esc_list = pa.esc_list; ca.esc_cap = pa.esc_cap; ca.esc_counters = pa.esc_counters;
        ca.esc_total = reinterpret_cast<unsigned*>(flag + 8);
        ca.brare_list = brare_list; ca.brare_cap = brare_cap(Npad, Kpad); ca.brare_count = reinterpret_cast<unsigned*>(flag + 13);
        ca.Aw = Aw; ca.Bw = Bw; ca.Mpad = Mpad; ca.Npad = Npad; ca.N = N; ca.Mrows = Mrows; ca.lut16 = d_lut; ca.c8_bad = c8src.bad;
        ca.tb_eff = flag + 12; ca.tmax = flag + 4; ca.NT = geom.NT; ca.M = M; ca.sidx = geom.sidx; ca.nt_magic = nt_magic;
        ca.zero_word = 2u * (uint32_t)geom.zeroA_off; ca.D = c8_D;
        ca.esc_limit = std::min<int64_t>((int64_t)pa.esc_cap * kEscLists, Mrows * K / 1024); ca.Kpad = (int)Kpad; ca.t_bias = geom.RA + M + 2;
        c8_correct_kernel<<<(unsigned)(8 * sms), 256, 0, st>>>(ca);
        B200_LAUNCH_CHECK("c8_correct_kernel");
    }

    LutGemmArgs la;
    la.Aw = Aw; la.Bw = Bw; la.lut = d_lut; la.C = C; la.ldc = ldc; la.Mrows = Mrows; la.N = N;
    la.Mpad = Mpad; la.Npad = Npad; la.Kpad = (int)Kpad; la.lut_bytes = lut_bytes;
    la.bias = bias; la.M = M;
    la.zero_word = 2u * (uint32_t)geom.zeroA_off;
    // k-loop choice (measured, 4096^3 E4M3: branchy 5.2 / 7.3 T MAC/s at 0 % / 50 % zeros, predicated 6.4 / 6.9)
    { const char* e = getenv("B200_LUT_SPARSE"); la.sparse_a = e ? atoi(e) : -1; }
    la.zero_count = reinterpret_cast<const unsigned long long*>(flag + 2);
    la.offgrid = flag;
    la.qamaa = qop ? 1 : 0;
    if (qop) la.qop = *qop; else la.qop = QamaaOp{};
    la.partial = partial; la.inexact = flag + 6; la.tmax = flag + 4; la.t_bias = geom.RA + M + 2;
    la.cp = cp;
    la.ep = ep;
    la.c8 = c8 ? 1 : 0;
    la.lut8 = c8src.lut8; la.lut8_bytes = c8src.lut8_bytes;
    la.esc_total = reinterpret_cast<const unsigned*>(flag + 8);
    la.esc_limit = std::min<int64_t>((int64_t)pa.esc_cap * kEscLists, Mrows * K / 1024);   // more: the correction pass eats the gain
    la.brare_count = reinterpret_cast<const unsigned*>(flag + 13);
    la.c8_D = c8_D;
    la.c8_ran = flag + 11;
    la.use_tma = 0;
    memset(&la.mapA, 0, sizeof(la.mapA));
    memset(&la.mapB, 0, sizeof(la.mapB));
    // one (tile, split) configuration: the GEMM, and for split-K the ordered reduction plus the gated unsplit kernel
    // TMA operand staging for the GEMM-form planes (B200_LUT_TMA=0: the cp.async ring, for before / after measurements)
    const bool want_tma = !gather && Mpad < ((int64_t)1 << 31) && Npad < ((int64_t)1 << 31) &&
                          !(getenv("B200_LUT_TMA") && atoi(getenv("B200_LUT_TMA")) == 0);
    auto set_maps = [&](LutGemmArgs& l, const LutTileCfg* cfg) -> int {
        l.use_tma = 0;
        if (!want_tma || cfg->BM() > 256 || cfg->BN() > 256) return B200_OK;
        if (int rc = lut_plane_map(&l.mapA, Aw, Mpad, Kpad, cfg->BM(), cfg->kt)) return rc;
        if (int rc = lut_plane_map(&l.mapB, Bw, Npad, Kpad, cfg->BN(), cfg->kt)) return rc;
        l.use_tma = 1;
        return B200_OK;
    };
    auto launch = [&](const LutTileCfg* cfg, int splits) -> int {
        LutGemmArgs l = la;
        if (l.c8 && (!cfg->kernel8() || !cfg->sk_kernel8())) l.c8 = 0;
        if (int rc = set_maps(l, cfg)) return rc;
        l.splits = splits; l.gate = nullptr;
        l.sk_ctas = 0; l.sk_tiles_n = 0; l.sk_tiles = 0;
        const dim3 tgrid((unsigned)ceil_div(Mrows, cfg->BM()), (unsigned)ceil_div(N, cfg->BN()), (unsigned)std::max(splits, 1));
        if (splits == 0) {   // stream-K
            const int64_t tiles = (int64_t)tgrid.x * tgrid.y, nk_all = Kpad / cfg->kt;
            const int64_t G = std::min<int64_t>((int64_t)sms * resident_ctas(*cfg, lut_bytes), tiles * nk_all);
            l.sk_ctas = (int)G; l.sk_tiles_n = (int)tgrid.y; l.sk_tiles = tiles;
            l.splits = (int)(2 + nk_all / std::max<int64_t>(1, tiles * nk_all / G));   // segments per tile, upper bound
            (gather ? cfg->sk_conv_kernel : cfg->sk_kernel)<<<dim3((unsigned)G), 32 * cfg->warps, cfg->smem(lut_bytes), st>>>(l);
            B200_LAUNCH_CHECK("approx_gemm_lut_kernel(stream-K)");
            if (l.c8) {   // the byte-cell twin: same units, same slots; one of the two exits at once (c8_selected)
                cfg->sk_kernel8()<<<dim3((unsigned)G), 32 * cfg->warps, cfg->smem(l.lut8_bytes), st>>>(l);
                B200_LAUNCH_CHECK("approx_gemm_lut_kernel(stream-K, byte cells)");
            }
            streamk_fixup_kernel<<<dim3((unsigned)tiles, 16), 256, 0, st>>>(partial, (int)G, tiles, (int)tgrid.y, (int)nk_all, cfg->BM(),
                                                                  cfg->BN(), C, ldc, Mrows, N, bias, M, l.qamaa, l.qop,
                                                                  flag + 6, flag, ep);
            B200_LAUNCH_CHECK("streamk_fixup_kernel");
        } else {
            (gather ? cfg->conv_kernel : cfg->kernel)<<<tgrid, 32 * cfg->warps, cfg->smem(lut_bytes), st>>>(l);
            if (l.c8) {   // the byte-cell twin on the same planes; one of the two exits at once (c8_selected)
                B200_LAUNCH_CHECK("approx_gemm_lut_kernel");
                cfg->kernel8()<<<tgrid, 32 * cfg->warps, cfg->smem(l.lut8_bytes), st>>>(l);
            }
        }
        if (splits > 1) {
            B200_LAUNCH_CHECK("approx_gemm_lut_kernel(split-K)");
            if (ep.on && ep.HW > 0)
                splitk_reduce_nchw_kernel<<<dim3((unsigned)ceil_div(Mrows, 32), (unsigned)ceil_div(N, 32)), 256, 0, st>>>(
                    partial, splits, C, Mrows, N, bias, M, l.qamaa, l.qop, flag + 6, flag, ep);
            else
                splitk_reduce_kernel<<<(unsigned)std::min<int64_t>(ceil_div(Mrows * N, 256), 64 * sms), 256, 0, st>>>(
                    partial, splits, C, ldc, Mrows, N, bias, M, l.qamaa, l.qop, flag + 6, flag, ep);
            B200_LAUNCH_CHECK("splitk_reduce_kernel");
        }
        if (splits != 1) {
            // runs only if the exactness certificate failed
            int S1 = 1;
            const LutTileCfg* t1 = choose_lut_tile(Mrows, N, Kpad, lut_bytes, geom.sidx, sms, 1, &S1);
            l.splits = 1; l.gate = flag + 6; l.sk_ctas = 0;
            if (int rc = set_maps(l, t1)) return rc;
            (gather ? t1->conv_kernel : t1->kernel)<<<dim3((unsigned)ceil_div(Mrows, t1->BM()), (unsigned)ceil_div(N, t1->BN())), 32 * t1->warps,
                                                     t1->smem(lut_bytes), st>>>(l);
        }
        B200_LAUNCH_CHECK("approx_gemm_lut_kernel");
        return B200_OK;
    };
    if (partial && getenv("B200_LUT_STREAMK") && atoi(getenv("B200_LUT_STREAMK")) != 0 && partial) {   // experiments / tests
        const int64_t tiles = ceil_div(Mrows, tc->BM()) * ceil_div(N, tc->BN());
        if ((size_t)(2 * (int64_t)sms * resident_ctas(*tc, lut_bytes)) * tc->BM() * tc->BN() * 4 <= kStreamKBytes && tiles > 0) {
            S = 0;
            forced = true;
        }
    }
    // measured choice (see TuneKey above)
    if (autotune_on() && !forced && cands.size() > 1 && Mrows * N * Kpad >= kTuneMinMacs) {
        const TuneKey key{Mrows, N, Kpad, lut_bytes, geom.sidx, gather ? 1 : 0, max_s, qop ? 1 : 0};
        bool have = false;
        {
            std::lock_guard<std::mutex> lk(g_tuned_mu);
            auto it = g_tuned.find(key);
            if (it != g_tuned.end()) { tc = it->second.first; S = it->second.second; have = true; }
        }
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (!have && cudaStreamIsCapturing(st, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone) {
            // the model's best regular configurations (within 1.6x of its favourite) plus the two best stream-K ones
            auto by_cost = [](const LutCand& x, const LutCand& y) { return x.cost < y.cost; };
            std::vector<LutCand> reg, sk;
            for (const LutCand& c : cands) (c.S == 0 ? sk : reg).push_back(c);
            std::sort(reg.begin(), reg.end(), by_cost);
            std::sort(sk.begin(), sk.end(), by_cost);
            cands.clear();
            for (size_t i = 0; i < reg.size() && (int)i < kTuneCandidates - 1 && reg[i].cost <= 1.6 * reg[0].cost; ++i) cands.push_back(reg[i]);
            for (size_t i = 0; i < sk.size() && i < 2; ++i) cands.push_back(sk[i]);
            cudaEvent_t e0, e1;
            B200_CUDA_TRY(cudaEventCreate(&e0));
            B200_CUDA_TRY(cudaEventCreate(&e1));
            int rc = B200_OK;
            auto time_cand = [&](const LutCand& c, float* ms) {
                for (int rep = 0; rep < 4 && rc == B200_OK; ++rep) {   // one warm run, three timed
                    if (rep == 1) cudaEventRecord(e0, st);
                    cudaMemsetAsync(flag + 6, 0, 4, st);
                    rc = launch(c.cfg, c.S);
                }
                if (rc != B200_OK) return;
                cudaEventRecord(e1, st);
                if (cudaEventSynchronize(e1) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "autotune"); return; }
                cudaEventElapsedTime(ms, e0, e1);
                g_tune_runs.fetch_add(1, std::memory_order_relaxed);
            };
            // The model's favourite is timed first AND last (clocks ramp up during the first runs of a process);
            // a challenger has to beat its better time by 1.5 %: timing noise must not displace the model's choice.
            std::vector<float> ms(cands.size(), 0.0f);
            for (size_t i = 0; i < cands.size() && rc == B200_OK; ++i) time_cand(cands[i], &ms[i]);
            if (rc == B200_OK && cands.size() > 1) {
                float again = 0.0f;
                time_cand(cands[0], &again);
                if (rc == B200_OK && again < ms[0]) ms[0] = again;
            }
            if (rc == B200_OK) {
                size_t best = 0;
                for (size_t i = 1; i < cands.size(); ++i)
                    if (ms[i] < 0.985f * ms[0] && (best == 0 || ms[i] < ms[best])) best = i;
                tc = cands[best].cfg;
                S = cands[best].S;
            }
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            if (rc != B200_OK) return rc;
            cudaMemsetAsync(flag + 6, 0, 4, st);
            std::lock_guard<std::mutex> lk(g_tuned_mu);
            g_tuned[key] = std::make_pair(tc, S);
        }
    }
    prof_record(true, st);
    if (int rc = launch(tc, S)) return rc;
    g_last_splits = S;
    prof_record(false, st);
    return B200_OK;
}


}  // namespace b200

using namespace b200;

extern "C" int b200_set_autotune(int on)
{
    const int prev = autotune_on() ? 1 : 0;
    g_autotune.store(on ? 1 : 0, std::memory_order_relaxed);
    return prev;
}

extern "C" uint64_t b200_autotune_runs(void) { return g_tune_runs.load(std::memory_order_relaxed); }

extern "C" size_t b200_approx_gemm_prepared_b_bytes(int64_t K, int64_t N) { return prepared_b_bytes(K, N); }

extern "C" size_t b200_approx_gemm_workspace_bytes(int64_t Mrows, int64_t K, int64_t N)
{
    if (Mrows < 0 || K < 0 || N < 0) return 0;
    const int64_t Mpad = round_up(Mrows > 0 ? Mrows : 1, kPadM), Npad = round_up(N > 0 ? N : 1, kPadN);
    const int64_t Kpad = round_up(K > 0 ? K : 1, kPadK);
    return kFlagBytes + (size_t)Kpad * (size_t)(Mpad + Npad) * 4 + c8_extra_bytes(Mrows, N, Mpad, Npad, Kpad) + splitk_partial_bytes(Mrows, N);
}

size_t b200::conv_gather_workspace_bytes(int64_t nB, int64_t Cg, int64_t H, int64_t W, int64_t Kg)
{
    return conv_plane_bytes(nB * Cg * H * W, round_up(Kg > 0 ? Kg : 1, kPadK));
}

// Shared by the GEMM and conv entry points (conv.cu): `A` is either a row-major matrix or an
// implicit-im2col view of an NCHW tensor.
int b200::approx_gemm_impl(const ASrc& A, const float* B, int64_t sb_k, int64_t sb_n, float* C, int64_t ldc,
                           int64_t Mrows, int64_t K, int64_t N, const BiasSrc& bias, const b200_approx_params* p,
                           void* workspace, size_t workspace_bytes, void* stream, const Epilogue* epp)
{
    NvtxRange nvtx_range("b200_approx_gemm");
    const Epilogue ep = epp ? *epp : Epilogue{};
    if (int rc = check_device()) return rc;
    if (!p || Mrows < 0 || K < 0 || N < 0) { set_error("b200_approx_gemm: bad argument"); return B200_ERR_ARGUMENT; }
    const int E = p->expo_width, M = p->mant_width;
    const bool v5 = p->ofuf_mode == 1;
    const signed char* host_table = nullptr;
    if (p->custom_table) {
        if (E < 1 || M < 1 || M > 5 || E + M > 7) { set_error("Invalid combination of expo_width and mant_width"); return B200_ERR_INVALID_FORMAT; }
        host_table = intern_table(reinterpret_cast<const signed char*>(p->custom_table), M);
    } else if (int rc = v5 ? select_table_v5(E, M, p->withComp, p->dnsmp_factor, &host_table)
                           : select_table(E, M, p->withComp, p->dnsmp_factor, p->allow_unpinned_format, &host_table)) {
        return rc;
    }
    if (p->ofuf_mode > 1) { set_error("b200_approx_gemm: unknown ofuf_mode"); return B200_ERR_ARGUMENT; }
    if (Mrows == 0 || N == 0) return B200_OK;
    if ((K > 0 && (!A.p || !B)) || !C || (!bias.iB && !bias.fB)) { set_error("b200_approx_gemm: NULL operand"); return B200_ERR_ARGUMENT; }
    cudaStream_t st = (cudaStream_t)stream;
    const signed char* dev_table = nullptr;
    if (int rc = device_table(host_table, M, &dev_table)) return rc;

    ApproxPairOp op;
    op.bias = bias;
    op.cfg.E = E; op.cfg.M = M;
    op.cfg.with_approx = p->with_approx != 0 && dev_table != nullptr;   // zero table == no correction
    op.cfg.s2nn2s = p->with_s2nn2s_opt != 0;
    op.cfg.clip_OF = p->golden_clip_OF != 0;
    op.cfg.qbma = p->quant_btw_mult_accu != 0;
    op.cfg.quirk = p->single_column_quirk != 0;
    op.T = op.cfg.with_approx ? dev_table : nullptr;
    V5Flags vf{v5 ? 1 : 0, p->sim_hw_add_OFUF != 0, p->with_OF_opt != 0, p->with_UF_opt != 0};
    if (v5) {   // the v5 GEMM knows none of approx_v9's refinements and has no tensor-bias branch
        op.cfg.v5 = true;
        op.cfg.sim_ofuf = vf.sim != 0; op.cfg.of_opt = vf.of_opt != 0; op.cfg.uf_opt = vf.uf_opt != 0;
        op.cfg.quirk = false;
        op.T = dev_table;
    }
    const BSrc bsrc{B, sb_k, sb_n};
    const dim3 ggrid = pair_gemm_grid(Mrows, N);

    const bool lut_mode = !p->force_generic && K > 0 &&
                          (v5 || (p->with_s2nn2s_opt && p->quant_btw_mult_accu && !p->single_column_quirk));
    const LutEntry* lut = nullptr;
    if (lut_mode) {
        if (int rc = v5 ? get_lut(E, M, host_table, dev_table, 0, 0, &lut, false, &vf)
                        : get_lut(E, M, p->with_approx ? host_table : nullptr, p->with_approx ? dev_table : nullptr,
                                  p->with_approx != 0, p->golden_clip_OF != 0, &lut))
            return rc;
    }
    if (!lut || lut->flags.sign_quirk || lut->flags.unrepresentable) {
        pair_gemm_kernel<ApproxPairOp><<<ggrid, 256, 0, st>>>(A, bsrc, C, ldc, Mrows, K, N, op, nullptr, 0, ep);
        B200_LAUNCH_CHECK("pair_gemm_kernel<ApproxPairOp>");
        set_path(v5 ? "generic-v5" : "generic");
        return B200_OK;
    }

    const int64_t Mpad = round_up(Mrows, kPadM), Npad = round_up(N, kPadN), Kpad = round_up(K, kPadK);
    // flag block, both decoded planes, the byte-cell variant's escape lists and correction sums
    const size_t planes = kFlagBytes + (size_t)Kpad * (size_t)(Mpad + Npad) * 4;
    const size_t words = planes + c8_extra_bytes(Mrows, N, Mpad, Npad, Kpad), need = words;
    if (!workspace || workspace_bytes < need) {
        set_error("b200_approx_gemm: workspace too small (see b200_approx_gemm_workspace_bytes)");
        return B200_ERR_WORKSPACE;
    }
    int* flag = reinterpret_cast<int*>(workspace);
    uint32_t* Aw = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(workspace) + kFlagBytes);
    uint32_t* Bw = Aw + (size_t)Kpad * Mpad;
    if (!(p->prepared_b && p->prepared_b_ready))   // (the restore kernel of a prepared B clears the block itself)
        B200_CUDA_TRY(cudaMemsetAsync(flag, 0, kFlagClearBytes, st));
    C8Src c8src;
    if (lut->d_lut8 && !p->force_table16) {
        c8src.lut8 = lut->d_lut8; c8src.lut8_bytes = lut->lut8_bytes; c8src.bad = lut->d_bad;
        c8src.extra = reinterpret_cast<uint8_t*>(workspace) + planes;
    }

    // split-K partial sums: only if the caller's workspace has the room b200_approx_gemm_workspace_bytes asks for
    // (not in the v5 flavour: without sim_hw_add_OFUF its cells need not be integers of u, and the exactness
    // certificate below is stated for approx_v9's table)
    const size_t pbytes = splitk_partial_bytes(Mrows, N);
    float* partial = (!v5 && pbytes && workspace_bytes >= words + pbytes) ? reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + words) : nullptr;
    // whatever the caller provides beyond the GEMM-form workspace holds the decoded conv plane
    void* conv_ws = workspace_bytes > words + pbytes ? reinterpret_cast<char*>(workspace) + words + pbytes : nullptr;
    const size_t conv_ws_bytes = conv_ws ? workspace_bytes - words - pbytes : 0;
    PreparedB prep;
    if (p->prepared_b) {
        if (p->prepared_b_bytes < prepared_b_bytes(K, N) || ((uintptr_t)p->prepared_b & 15) != 0) {
            set_error("b200_approx_gemm: prepared_b too small or misaligned (see b200_approx_gemm_prepared_b_bytes)");
            return B200_ERR_WORKSPACE;
        }
        prep.base = reinterpret_cast<uint8_t*>(p->prepared_b);
        prep.ready = p->prepared_b_ready != 0;
    }
    if (int rc = run_lut_gemm(A, B, sb_k, sb_n, C, ldc, Mrows, K, N, bias, lut->geom, lut->d_lut, lut->lut_bytes, flag,
                              Aw, Bw, Mpad, Npad, Kpad, nullptr, partial, conv_ws, conv_ws_bytes, st, ep, c8src, prep))
        return rc;
    // runs only if the predecode found an off-grid operand
    pair_gemm_kernel<ApproxPairOp><<<ggrid, 256, 0, st>>>(A, bsrc, C, ldc, Mrows, K, N, op, flag, 1, ep);
    B200_LAUNCH_CHECK("pair_gemm_kernel<ApproxPairOp>(conditional)");
    if (v5) { set_path("lut-v5"); return B200_OK; }
    set_path(lut->geom.sidx ? "lut-signed" : (g_last_splits > 1 ? "lut-splitk" : (g_last_splits == 0 ? "lut-streamk" : "lut")));
    return B200_OK;
}

// quantize_after_mult_and_add on the table kernel.  `hint` carries the operand exponent biases (the
// fake-quantizers' custom_bias tensors): they only fix the table's index range -- an operand outside the
// grid they describe raises the device flag and the arithmetic kernel recomputes the output.
constexpr size_t kQamaaTable16Bytes = 128 * 1024, kQamaaTableBytes = 192 * 1024;   // 16-bit table + byte-cell companion

extern "C" size_t b200_qamaa_gemm_workspace_bytes(int64_t Mrows, int64_t K, int64_t N)
{
    const size_t base = b200_approx_gemm_workspace_bytes(Mrows, K, N);
    return base ? base + kQamaaTableBytes : 0;
}

int b200::qamaa_gemm_impl(const ASrc& A, const float* B, int64_t sb_k, int64_t sb_n, float* C, int64_t ldc,
                          int64_t Mrows, int64_t K, int64_t N, const float* maxval_dev, int n_bits,
                          float mantissa_bits, int sign_bits, const BiasSrc* hint, void* workspace,
                          size_t workspace_bytes, void* stream)
{
    NvtxRange nvtx_range("b200_qamaa_gemm");
    cudaStream_t st = (cudaStream_t)stream;
    QamaaOp op;
    op.maxval_dev = maxval_dev;
    op.n_bits = n_bits;
    op.sign_bits = sign_bits;
    op.mantissa_bits = mantissa_bits;
    const BSrc bsrc{B, sb_k, sb_n};
    const dim3 ggrid = pair_gemm_grid(Mrows, N);
    const int M = (int)rintf(mantissa_bits), E = n_bits - sign_bits - M;
    const bool lut_mode = hint && n_bits == 8 && sign_bits == 1 && M >= 2 && M <= 5 && E >= 2 && K > 0 && workspace &&
                          workspace_bytes >= b200_qamaa_gemm_workspace_bytes(Mrows, K, N) && !getenv("B200_QAMAA_GENERIC");
    if (!lut_mode) {
        pair_gemm_kernel<QamaaOp><<<ggrid, 256, 0, st>>>(A, bsrc, C, ldc, Mrows, K, N, op, nullptr, 0);
        B200_LAUNCH_CHECK("pair_gemm_kernel<QamaaOp>");
        set_path("qamaa-generic");
        return B200_OK;
    }
    const LutGeom geom = make_lut_geom(E, M);
    const int lut_bytes = (int)round_up((int64_t)geom.total_entries * 2, 16);
    const int lut8_bytes = (int)round_up((int64_t)geom.total_entries, 16);
    if ((size_t)lut_bytes > kQamaaTable16Bytes || (size_t)lut8_bytes > kQamaaTableBytes - kQamaaTable16Bytes) {
        set_error("b200_qamaa_gemm: table does not fit");
        return B200_ERR_ARGUMENT;
    }
    const int64_t Mpad = round_up(Mrows, kPadM), Npad = round_up(N, kPadN), Kpad = round_up(K, kPadK);
    int* flag = reinterpret_cast<int*>(workspace);
    float* bias_R = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + 64);
    uint32_t* Aw = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(workspace) + kFlagBytes);
    uint32_t* Bw = Aw + (size_t)Kpad * Mpad;
    uint16_t* d_lut = reinterpret_cast<uint16_t*>(Bw + (size_t)Kpad * Npad);
    B200_CUDA_TRY(cudaMemsetAsync(flag, 0, kFlagClearBytes, st));
    lut_build_qamaa_kernel<<<(int)ceil_div(lut_bytes / 2, 256), 256, 0, st>>>(geom, op, d_lut, bias_R, flag);
    B200_LAUNCH_CHECK("lut_build_qamaa_kernel");
    C8Src c8src;
    uint8_t* after_tables = reinterpret_cast<uint8_t*>(d_lut) + kQamaaTableBytes;
    if (c8_format(M)) {   // byte-cell companion of this call's table, its "bad row" word in the flag block
        int8_t* d_lut8 = reinterpret_cast<int8_t*>(reinterpret_cast<char*>(d_lut) + kQamaaTable16Bytes);
        lut8_build_kernel<<<(int)ceil_div(lut8_bytes, 256), 256, 0, st>>>(geom, d_lut, d_lut8, lut8_bytes, flag + 10);
        B200_LAUNCH_CHECK("lut8_build_kernel");
        c8src.lut8 = d_lut8; c8src.lut8_bytes = lut8_bytes; c8src.bad = flag + 10; c8src.extra = after_tables;
    }
    BiasSrc bias = *hint;
    bias.fR = bias_R;                                   // result bias = the quantizer's, computed on the device
    float* partial = splitk_partial_bytes(Mrows, N) ? reinterpret_cast<float*>(after_tables + c8_extra_bytes(Mrows, N, Mpad, Npad, Kpad)) : nullptr;
    const size_t base_bytes = b200_qamaa_gemm_workspace_bytes(Mrows, K, N);
    void* conv_ws = workspace_bytes > base_bytes ? reinterpret_cast<char*>(workspace) + base_bytes : nullptr;
    const size_t conv_ws_bytes = conv_ws ? workspace_bytes - base_bytes : 0;
    if (int rc = run_lut_gemm(A, B, sb_k, sb_n, C, ldc, Mrows, K, N, bias, geom, d_lut, lut_bytes, flag, Aw, Bw, Mpad,
                              Npad, Kpad, &op, partial, conv_ws, conv_ws_bytes, st, Epilogue{}, c8src))
        return rc;
    pair_gemm_kernel<QamaaOp><<<ggrid, 256, 0, st>>>(A, bsrc, C, ldc, Mrows, K, N, op, flag, 1);
    B200_LAUNCH_CHECK("pair_gemm_kernel<QamaaOp>(conditional)");
    set_path(g_last_splits > 1 ? "qamaa-lut-splitk" : (g_last_splits == 0 ? "qamaa-lut-streamk" : "qamaa-lut"));
    return B200_OK;
}

extern "C" int b200_approx_gemm(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, float* C,
                                int64_t ldc, int64_t Mrows, int64_t K, int64_t N, int bA, const int32_t* bB, int bR,
                                const b200_approx_params* p, void* workspace, size_t workspace_bytes, void* stream)
{
    BiasSrc b{bA, bR, nullptr, nullptr, bB, nullptr, 1};
    return approx_gemm_impl(make_gemm_asrc(A, lda), B, sb_k, sb_n, C, ldc, Mrows, K, N, b, p, workspace, workspace_bytes, stream);
}

// Sets the fused-input-quantizer fields of an operand view; B200_OK, or an error for a bad descriptor.
int b200::apply_act_quant(ASrc* a, const b200_act_quant* aq)
{
    if (!aq) return B200_OK;
    if (!aq->maxval_dev || !aq->bias_dev || aq->n_bits < 2 || aq->sign_bits < 0 || aq->sign_bits > 1) {
        set_error("b200_act_quant: bad descriptor");
        return B200_ERR_ARGUMENT;
    }
    a->fq_maxval = aq->maxval_dev;
    a->fq_bias = aq->bias_dev;
    a->fq_Mf = fminf(fmaxf(rintf(aq->mantissa_bits), 1.0f), (float)(aq->n_bits - aq->sign_bits));
    a->fq_sign = aq->sign_bits;
    return B200_OK;
}

extern "C" int b200_approx_gemm_fbias_fq(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n,
                                         float* C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bA_dev,
                                         const float* bB_dev, int64_t bB_stride, const float* bR_dev,
                                         const b200_approx_params* p, const b200_act_quant* aq, void* workspace,
                                         size_t workspace_bytes, void* stream)
{
    if (!p) { set_error("b200_approx_gemm_fbias_fq: params is NULL"); return B200_ERR_ARGUMENT; }
    const int dflt = 1 << (p->expo_width > 0 ? p->expo_width - 1 : 0);
    BiasSrc b{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev, bB_stride};
    ASrc a = make_gemm_asrc(A, lda);
    if (int rc = apply_act_quant(&a, aq)) return rc;
    return approx_gemm_impl(a, B, sb_k, sb_n, C, ldc, Mrows, K, N, b, p, workspace, workspace_bytes, stream);
}

int b200::make_epilogue(const b200_epilogue* e, Epilogue* out)
{
    *out = Epilogue{};
    if (!e) return B200_OK;
    if (e->act < 0 || e->act > 2) { set_error("b200_epilogue: act must be 0 (none), 1 (ReLU) or 2 (ReLU6)"); return B200_ERR_ARGUMENT; }
    if (e->bn_mean && !e->bn_var) { set_error("b200_epilogue: bn_mean without bn_var"); return B200_ERR_ARGUMENT; }
    out->on = 1;
    out->bias = e->bias;
    out->mean = e->bn_mean; out->var = e->bn_var; out->gamma = e->bn_gamma; out->beta = e->bn_beta;
    out->eps = e->bn_eps;
    out->act = e->act;
    return B200_OK;
}

extern "C" int b200_approx_gemm_fbias_ep(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n,
                                         float* C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bA_dev,
                                         const float* bB_dev, int64_t bB_stride, const float* bR_dev,
                                         const b200_approx_params* p, const b200_act_quant* aq, const b200_epilogue* ep,
                                         void* workspace, size_t workspace_bytes, void* stream)
{
    if (!p) { set_error("b200_approx_gemm_fbias_ep: params is NULL"); return B200_ERR_ARGUMENT; }
    const int dflt = 1 << (p->expo_width > 0 ? p->expo_width - 1 : 0);
    BiasSrc b{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev, bB_stride};
    ASrc a = make_gemm_asrc(A, lda);
    if (int rc = apply_act_quant(&a, aq)) return rc;
    Epilogue e;
    if (int rc = make_epilogue(ep, &e)) return rc;
    e.Ctot = (int)N;    // row-major store: channel == column
    return approx_gemm_impl(a, B, sb_k, sb_n, C, ldc, Mrows, K, N, b, p, workspace, workspace_bytes, stream, ep ? &e : nullptr);
}

extern "C" int b200_approx_gemm_fbias(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n,
                                      float* C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bA_dev,
                                      const float* bB_dev, int64_t bB_stride, const float* bR_dev,
                                      const b200_approx_params* p, void* workspace, size_t workspace_bytes,
                                      void* stream)
{
    if (!p) { set_error("b200_approx_gemm_fbias: params is NULL"); return B200_ERR_ARGUMENT; }
    const int dflt = 1 << (p->expo_width > 0 ? p->expo_width - 1 : 0);   // approx_calculation.py:938-939
    BiasSrc b{dflt, dflt, bA_dev, bR_dev, nullptr, bB_dev, bB_stride};
    return approx_gemm_impl(make_gemm_asrc(A, lda), B, sb_k, sb_n, C, ldc, Mrows, K, N, b, p, workspace, workspace_bytes, stream);
}

extern "C" int b200_profile_begin(void)
{
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& e : g_prof_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
    g_prof_events.clear();
    g_prof_on = true;
    return B200_OK;
}

extern "C" int b200_profile_end(double* total_ms, int* launches)
{
    std::lock_guard<std::mutex> lk(g_prof_mu);
    double tot = 0.0;
    int n = 0;
    for (auto& e : g_prof_events) {
        float ms = 0.0f;
        if (cudaEventSynchronize(e.second) == cudaSuccess && cudaEventElapsedTime(&ms, e.first, e.second) == cudaSuccess) {
            tot += ms;
            ++n;
        }
        cudaEventDestroy(e.first);
        cudaEventDestroy(e.second);
    }
    g_prof_events.clear();
    g_prof_on = false;
    if (total_ms) *total_ms = tot;
    if (launches) *launches = n;
    return B200_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of b200_qamaa_gemm_fbias (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops).
// The scratch the table kernel needs (flags, both decoded operand planes, the pair table rebuilt from `maxval`,
// the split-K partial sums) lives in ONE file-scope device array instead of a caller workspace.  Every GPU that
// loads the library holds these B200_STATIC_SCRATCH_BYTES until the process exits.
//   * In host code `g_static_scratch` is only a stand-in for the symbol: its device address is looked up once
//     per device (with that device current) through cudaGetSymbolAddress and that pointer goes to the kernels.
//   * All calls on a device share the array and it keeps the previous call's contents: the flag words are
//     cleared at the start of every call (qamaa_gemm_impl: cudaMemsetAsync of the flag words), planes, table and
//     partial sums are fully rewritten before they are read.  Calls that overlap on the GPU (two streams, two
//     host threads) RACE on it -- serialise them.
//   * No cudaMalloc / cudaFree / allocating library routine on this path.
// ------------------------------------------------------------------------------------------------
constexpr size_t kStaticScratchBytes = (size_t)64 << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

namespace {
std::mutex g_static_mu;
void* g_static_ptr[b200::PerDeviceOnce::kMaxDev] = {};

// device address of g_static_scratch on the CURRENT device (looked up once per device)
int static_scratch_ptr(void** out)
{
    int dev = 0;
    B200_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= b200::PerDeviceOnce::kMaxDev) { set_error("device index out of range"); return B200_ERR_ARGUMENT; }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_scratch)");
        if (!p || ((uintptr_t)p & 255) != 0) { set_error("b200_qamaa_gemm_fbias_static: bad scratch address"); return B200_ERR_CUDA; }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return B200_OK;
}
}  // namespace

extern "C" size_t b200_static_scratch_bytes(void) { return kStaticScratchBytes; }

extern "C" int b200_qamaa_gemm_fbias_static(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n,
                                            float* C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N,
                                            const float* maxval_dev, int n_bits, float mantissa_bits, int sign_bits,
                                            const float* bA_dev, const float* bB_dev, int64_t bB_stride, void* stream)
{
    if (int rc = check_device()) return rc;
    if (Mrows < 0 || K < 0 || N < 0 || !maxval_dev || !bB_dev) {
        set_error("b200_qamaa_gemm_fbias_static: bad argument (the operand biases are required)");
        return B200_ERR_ARGUMENT;
    }
    if (Mrows == 0 || N == 0) return B200_OK;
    if (!A || !B || !C) { set_error("b200_qamaa_gemm_fbias_static: NULL operand"); return B200_ERR_ARGUMENT; }
    const size_t need = b200_qamaa_gemm_workspace_bytes(Mrows, K, N);
    if (need == 0 || need > kStaticScratchBytes) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "b200_qamaa_gemm_fbias_static: problem %lld x %lld x %lld needs %zu bytes of scratch, the static array holds "
                 "%zu (use b200_qamaa_gemm_fbias with a caller workspace)",
                 (long long)Mrows, (long long)K, (long long)N, need, kStaticScratchBytes);
        set_error(msg);
        return B200_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    if (int rc = static_scratch_ptr(&scratch)) return rc;
    const int E = n_bits - sign_bits - (int)rintf(mantissa_bits);
    const int dflt = 1 << (E > 0 ? E - 1 : 0);
    BiasSrc hint{dflt, dflt, bA_dev, nullptr, nullptr, bB_dev, bB_stride};
    // exactly the workspace size the allocating entry point is given: same tile / split-K choice, same bits
    return qamaa_gemm_impl(make_gemm_asrc(A, lda), B, sb_k, sb_n, C, ldc, Mrows, K, N, maxval_dev, n_bits, mantissa_bits,
                           sign_bits, &hint, scratch, need, stream);
}
