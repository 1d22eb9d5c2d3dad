// lut_gemm.cuh -- the table-driven GEMM kernel template (approx_gemm.cu instantiates the 16-bit-cell variants,
// approx_gemm_c8.cu the byte-cell ones) and the argument block both share.
#pragma once
#include <cuda.h>

#include <type_traits>

#include "common.cuh"
#include "epilogue.cuh"
#include "gemm_impl.cuh"
#include "lut.cuh"
#include "qamaa_op.cuh"

namespace b200 {

// Conv form, second variant: the group's input channels are decoded ONCE, in place of the kh*kw-times
// inflated plane above, into words Pw[b][c][h][w]; the GEMM's A-tile loader then does the im2col gather
// itself (4-byte cp.async from this L2-resident plane).  kinfo[k] = {word offset of tap k = (c, ky, kx)
// relative to the pixel's top-left tap, ky*dh | kx*dw << 16}; k >= K gets an always-out-of-range row offset.
struct ConvPlane {
    const uint32_t* Pw;       // [B][Cg][H][W] words
    const int2* kinfo;        // [Kpad]
    const uint32_t* zero_src; // one word: the A word of a == 0 (padding, rows / k beyond the matrix)
    int64_t img_words;        // Cg * H * W
    int H, W, Ho, Wo, sh, sw, ph, pw;
};

// ------------------------------------------------------------------------------------------------
// Table-driven GEMM.  Template <WARPS, RPW, CPL>: CTA tile (RPW*WARPS) x (32*CPL); one warp =
// RPW rows x 32*CPL columns; one lane = RPW rows x CPL consecutive columns (RPW*CPL = 64 or 32 fp32
// accumulators).  For a fixed k all lanes of a warp use the same RPW A words (broadcast LDS.128) and
// their own CPL B words (one conflict-free LDS.128 / LDS.64).  Operand tiles are staged through a
// 3-deep cp.async ring; the pair table lives in shared memory for the life of the CTA.
// Per MAC: IADD (table address) + LDS.U16 + FHFMA.BF16.
// ------------------------------------------------------------------------------------------------
constexpr int LUT_STAGES = 3;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned short lds_u16(uint32_t addr)
{
    unsigned short v;
    asm("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(addr));
    return v;
}
// acc += e * m with e, m bf16 and acc fp32 (PTX ISA 8.6 mixed-precision fma, SASS FHFMA.BF16):
// the product of two bf16 is exact in fp32 and the add rounds once -- identical to an fp32 add of
// the (exactly representable) table value.
__device__ __forceinline__ void fma_bf16(float& acc, unsigned short e, unsigned short m)
{
    asm("fma.rn.f32.bf16 %0, %1, %2, %0;" : "+f"(acc) : "h"(e), "h"(m));
}
// One A element against CPL B elements: table address add, 16-bit table load, fp32 += bf16 * bf16,
// skipped as a whole when a == 0.  The test is warp-uniform (all lanes of a warp hold the same A
// word), which `bra.uni` tells ptxas: a plain branch, no BSSY/BSYNC reconvergence pair.  Zero
// activations -- about half of a post-ReLU tensor -- issue no shared-memory lookups at all.
__device__ __forceinline__ void mac_row4(float& c0, float& c1, float& c2, float& c3, uint32_t aw, uint32_t zero_word,
                                         uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3, unsigned short m0,
                                         unsigned short m1, unsigned short m2, unsigned short m3)
{
    asm("{\n\t.reg .pred p;\n\t.reg .b16 e0, e1, e2, e3;\n\t.reg .u32 x0, x1, x2, x3;\n\t"
        "setp.ne.u32 p, %4, %5;\n\t"
        "add.u32 x0, %4, %6;\n\tadd.u32 x1, %4, %7;\n\tadd.u32 x2, %4, %8;\n\tadd.u32 x3, %4, %9;\n\t"
        "@p ld.shared.u16 x0, [x0];\n\t@p ld.shared.u16 x1, [x1];\n\t@p ld.shared.u16 x2, [x2];\n\t@p ld.shared.u16 x3, [x3];\n\t"
        "cvt.u16.u32 e0, x0;\n\tcvt.u16.u32 e1, x1;\n\tcvt.u16.u32 e2, x2;\n\tcvt.u16.u32 e3, x3;\n\t"
        "@p fma.rn.f32.bf16 %0, e0, %10, %0;\n\t@p fma.rn.f32.bf16 %1, e1, %11, %1;\n\t"
        "@p fma.rn.f32.bf16 %2, e2, %12, %2;\n\t@p fma.rn.f32.bf16 %3, e3, %13, %3;\n\t}"
        : "+f"(c0), "+f"(c1), "+f"(c2), "+f"(c3)
        : "r"(aw), "r"(zero_word), "r"(b0), "r"(b1), "r"(b2), "r"(b3), "h"(m0), "h"(m1), "h"(m2), "h"(m3));
}
__device__ __forceinline__ void mac_row2(float& c0, float& c1, uint32_t aw, uint32_t zero_word, uint32_t b0,
                                         uint32_t b1, unsigned short m0, unsigned short m1)
{
    asm("{\n\t.reg .pred p;\n\t.reg .b16 e0, e1;\n\t.reg .u32 x0, x1;\n\t"
        "setp.ne.u32 p, %2, %3;\n\t"
        "add.u32 x0, %2, %4;\n\tadd.u32 x1, %2, %5;\n\t"
        "@p ld.shared.u16 x0, [x0];\n\t@p ld.shared.u16 x1, [x1];\n\t"
        "cvt.u16.u32 e0, x0;\n\tcvt.u16.u32 e1, x1;\n\t"
        "@p fma.rn.f32.bf16 %0, e0, %6, %0;\n\t@p fma.rn.f32.bf16 %1, e1, %7, %1;\n\t}"
        : "+f"(c0), "+f"(c1)
        : "r"(aw), "r"(zero_word), "r"(b0), "r"(b1), "h"(m0), "h"(m1));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// TMA operand staging (GEMM form): the k-major planes Aw[Kpad][Mpad] / Bw[Kpad][Npad] are dense 2-D tensors, one
// cp.async.bulk.tensor.2d per operand and k tile lands a [KT][BM] / [KT][BN] box in shared memory and signals a
// per-stage mbarrier with the byte count; issued by ONE thread, no address arithmetic or LDGSTS in the other 255.
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "LUT_WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra LUT_WAIT_DONE;\n\t"
        "bra LUT_WAIT_LOOP;\n\t"
        "LUT_WAIT_DONE:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1) : "memory");
}

struct LutGemmArgs {
    const uint32_t* Aw;   // [Kpad][Mpad]
    const uint32_t* Bw;   // [Kpad][Npad]
    const uint16_t* lut;  // total_entries (padded to 16 B)
    float* C;
    int64_t ldc, Mrows, N, Mpad, Npad;
    int Kpad;
    int lut_bytes;        // multiple of 16
    BiasSrc bias;
    int M;                // unit u = 2^(1 - bR - M)
    uint32_t zero_word;   // the A word of a == 0
    int sparse_a;         // 1 / 0: force the branchy / predicated k loop; -1: decide from zero_count
    const unsigned long long* zero_count;   // zero activations counted by the predecode pass
    const int* offgrid;
    int qamaa;            // 1: epilogue applies the result quantizer to the sum (quantize_after_mult_and_add)
    QamaaOp qop;
    // split-K (grids that cannot fill the SMs): blockIdx.z owns a slice of the k tiles and stores its partial
    // sums, in units of u, to `partial`; splitk_reduce_kernel adds the slices in order.  That changes the
    // summation order, so it is only USED when every sum involved is provably exact in fp32 (see below);
    // otherwise *inexact is raised and the unsplit kernel (gate != null) recomputes the output.
    int splits;
    float* partial;       // [splits][Mrows][N]
    int* inexact;
    const int* tmax;      // [2]: largest table row index of A / of B seen by the predecode pass
    int t_bias;           // t = tA_idx + tB_idx - t_bias
    const int* gate;      // non-null: run only if *gate != 0
    ConvPlane cp;         // CONV instantiations: the A operand is gathered from a decoded NCHW plane
    // stream-K (sk_ctas > 0): a 1-D grid of sk_ctas persistent CTAs; the (tile, k-tile) work units, tiles in
    // column-fastest order, are dealt in equal contiguous ranges.  A CTA that covers a tile's whole K range
    // stores it directly; partial ranges go to `partial` as [cta][slot][BM][BN] (slot 0: the CTA's first
    // segment, slot 1: its last) and streamk_fixup_kernel adds a tile's segments in k order -- under the same
    // exactness certificate as split-K (`splits` = upper bound of segments per tile).
    int sk_ctas, sk_tiles_n;
    int64_t sk_tiles;
    Epilogue ep;          // fused +bias -> BN -> activation -> (NCHW) store, epilogue.cuh; ep.on == 0: plain store
    int use_tma;          // GEMM form only: operand tiles arrive by TMA (mapA / mapB) instead of the cp.async ring
    // byte-cell variant (C8 instantiations, approx_gemm_c8.cu): the same table with int8 cells in shared memory (half the
    // footprint a warp's lookups spread over => fewer bank conflicts) and integer accumulators.  Cells that do not fit
    // int8 are ZERO in that table; the pairs that meet one involve an "escape" A element (listed by the predecode pass),
    // and c8_correct_kernel has added their 16-bit cells into `c8_D` before this kernel starts from it.  Both variants
    // are enqueued on the same operand planes; each CTA evaluates c8_selected() and the wrong variant exits at once.
    int c8;                         // 1: a byte-cell launch accompanies this one
    const int8_t* lut8;             // total_entries int8 cells (padded to 16 B)
    int lut8_bytes;
    const unsigned* esc_total;      // [0] number of escape A elements, [1] list overflow (c8_correct_kernel)
    int64_t esc_limit;              // the gate closes above this many escapes
    const unsigned* brare_count;    // rare B elements (their pairs with non-escape A rows are in c8_D too)
    const int* c8_D;                // [Mrows][N] sums of the 16-bit cells the byte table dropped, in units of u
    int* c8_ran;                    // set to 1 by the byte-cell launch when it is the one that computes (tests, ops.table_variant)
    alignas(64) CUtensorMap mapA;   // Aw as a 2-D tensor {Mpad, Kpad}, box {BM, KT}
    alignas(64) CUtensorMap mapB;   // Bw as a 2-D tensor {Npad, Kpad}, box {BN, KT}
};

// Byte-cell variant or 16-bit variant?  Evaluated identically by every CTA of both launches (device-side flags only):
// the byte cells are used when (i) the integer sum is the fp32 k-ascending sum -- every |cell| < 2^dl with
// dl = t_max + M + 1 and Kpad * 2^dl <= 2^24, so every partial sum in ANY order is an integer below 2^24, exact in
// fp32 -- and (ii) the escapes fitted their lists and are few enough for the correction pass not to eat the gain.
__device__ __forceinline__ bool c8_selected(const LutGemmArgs& a)
{
    if (!a.c8) return false;
    const int dl = a.tmax[0] + a.tmax[1] - a.t_bias + a.M + 1;
    const int lgK = 32 - __clz((a.Kpad > 2 ? a.Kpad : 2) - 1);
    if (dl + lgK > 24) return false;
    return a.esc_total[1] == 0 && (int64_t)a.esc_total[0] <= a.esc_limit;
}
__device__ __forceinline__ int lds_s8(uint32_t addr)
{
    int v;
    asm("ld.shared.s8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
// byte-cell twins of mac_row4 / mac_row2: c += cell * (+-1) in int32; `ah` is the A word already halved
__device__ __forceinline__ void mac8_row4(int& c0, int& c1, int& c2, int& c3, uint32_t aw, uint32_t zero_word, uint32_t b0,
                                          uint32_t b1, uint32_t b2, uint32_t b3, int m0, int m1, int m2, int m3)
{
    asm("{\n\t.reg .pred p;\n\t.reg .u32 h;\n\t.reg .s32 x0, x1, x2, x3;\n\t"
        "setp.ne.u32 p, %4, %5;\n\tshr.u32 h, %4, 1;\n\t"
        "add.s32 x0, h, %6;\n\tadd.s32 x1, h, %7;\n\tadd.s32 x2, h, %8;\n\tadd.s32 x3, h, %9;\n\t"
        "@p ld.shared.s8 x0, [x0];\n\t@p ld.shared.s8 x1, [x1];\n\t@p ld.shared.s8 x2, [x2];\n\t@p ld.shared.s8 x3, [x3];\n\t"
        "@p mad.lo.s32 %0, x0, %10, %0;\n\t@p mad.lo.s32 %1, x1, %11, %1;\n\t"
        "@p mad.lo.s32 %2, x2, %12, %2;\n\t@p mad.lo.s32 %3, x3, %13, %3;\n\t}"
        : "+r"(c0), "+r"(c1), "+r"(c2), "+r"(c3)
        : "r"(aw), "r"(zero_word), "r"(b0), "r"(b1), "r"(b2), "r"(b3), "r"(m0), "r"(m1), "r"(m2), "r"(m3));
}
__device__ __forceinline__ void mac8_row2(int& c0, int& c1, uint32_t aw, uint32_t zero_word, uint32_t b0, uint32_t b1,
                                          int m0, int m1)
{
    asm("{\n\t.reg .pred p;\n\t.reg .u32 h;\n\t.reg .s32 x0, x1;\n\t"
        "setp.ne.u32 p, %2, %3;\n\tshr.u32 h, %2, 1;\n\t"
        "add.s32 x0, h, %4;\n\tadd.s32 x1, h, %5;\n\t"
        "@p ld.shared.s8 x0, [x0];\n\t@p ld.shared.s8 x1, [x1];\n\t"
        "@p mad.lo.s32 %0, x0, %6, %0;\n\t@p mad.lo.s32 %1, x1, %7, %1;\n\t}"
        : "+r"(c0), "+r"(c1)
        : "r"(aw), "r"(zero_word), "r"(b0), "r"(b1), "r"(m0), "r"(m1));
}

// KT: k-tile depth (8 where a large table leaves little shared memory for the operand ring).
// SIDX: signed-index table layout (lut.cuh): the B word is a full 32-bit byte offset, multiplier +1.0.
// CONV: implicit im2col in the A-tile loader (ConvPlane) instead of a pre-gathered Aw plane.
// SK: stream-K instantiation (persistent CTAs over (tile, k-tile) ranges, see LutGemmArgs::sk_ctas).
// C8: byte-cell variant (see LutGemmArgs::c8); GEMM-form planes only.
template <int WARPS, int RPW, int CPL, int KT, bool SIDX, bool CONV, bool SK, bool C8 = false>
__global__ void __launch_bounds__(32 * WARPS, (WARPS >= 16) ? 1 : ((RPW * CPL >= 64) ? 16 / WARPS : 24 / WARPS))
approx_gemm_lut_kernel(const __grid_constant__ LutGemmArgs a)
{
    constexpr int BM = RPW * WARPS, BN = 32 * CPL, ST = LUT_STAGES, NTHR = 32 * WARPS;
    static_assert(CPL == 4 || CPL == 2, "a lane owns 2 or 4 consecutive columns");
    static_assert(RPW % 4 == 0, "A words are read 4 at a time");
    if (*a.offgrid != 0) return;   // some operand is off the FP8 grid: the generic kernel takes over
    if (a.gate && *a.gate == 0) return;
    static_assert(!(C8 && CONV), "the byte-cell variant reads pre-gathered planes");
    if (c8_selected(a) != C8) return;
    if (C8 && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) *a.c8_ran = 1;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t full_bar[2 * ST];                // TMA: [s] "stage s has landed", [ST + s] "every warp has read stage s"
    const int tab_bytes = C8 ? a.lut8_bytes : a.lut_bytes;
    uint32_t* As = reinterpret_cast<uint32_t*>(smem + ((tab_bytes + 127) & ~127));   // [ST][KT][BM], 128-byte aligned (TMA)
    uint32_t* Bs = As + ST * KT * BM;                                 // [ST][KT][BN]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool tma = !CONV && a.use_tma != 0;
    int64_t m0 = (int64_t)blockIdx.x * BM, n0 = (int64_t)blockIdx.y * BN;

    // pair table -> shared memory
    {
        const uint4* src = C8 ? reinterpret_cast<const uint4*>(a.lut8) : reinterpret_cast<const uint4*>(a.lut);
        uint4* dst = reinterpret_cast<uint4*>(smem);
        for (int i = tid; i < tab_bytes / 16; i += NTHR) dst[i] = __ldg(src + i);
    }
    const uint32_t lut_base = smem_u32(smem);
    const uint32_t As_u = smem_u32(As), Bs_u = smem_u32(Bs);
    const uint32_t bar_u = smem_u32(full_bar);
    if (tma) {
        if (tid == 0) {
#pragma unroll
            for (int s = 0; s < ST; ++s) { mbar_init(bar_u + 8u * s, 1); mbar_init(bar_u + 8u * (ST + s), WARPS); }
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();   // nobody polls a barrier before it is initialised
    }

    // CONV: this thread gathers row (tid % BM) of every A tile, k = tid / BM + j * (NTHR / BM)
    static_assert(NTHR % BM == 0 && (KT * BM) % NTHR == 0, "A-tile gather mapping");
    const uint32_t* crow = nullptr;   // address of the row's top-left tap (may lie outside the image: never read)
    int ch0 = 0, cw0 = 0;
    auto conv_row_setup = [&]() {
        const int64_t r = m0 + tid % BM;
        const int wo = (int)(r % a.cp.Wo);
        const int64_t t2 = r / a.cp.Wo;
        const int ho = (int)(t2 % a.cp.Ho);
        const int64_t b = t2 / a.cp.Ho;
        ch0 = ho * a.cp.sh - a.cp.ph;
        cw0 = wo * a.cp.sw - a.cp.pw;
        if (r >= a.Mrows) ch0 = -(1 << 20);   // rows beyond the matrix: every tap out of range
        crow = a.cp.Pw + b * a.cp.img_words + (int64_t)ch0 * a.cp.W + cw0;
    };
    if (CONV) conv_row_setup();
    uint32_t greg[CONV ? KT * BM / NTHR : 1];   // CONV: this thread's gathered A words of the tile in flight
    auto gather_store = [&](int stage) {
        uint32_t* dst = As + (stage * KT + tid / BM) * BM + tid % BM;
#pragma unroll
        for (int j = 0; j < KT * BM / NTHR; ++j) dst[j * (NTHR / BM) * BM] = greg[j];
    };
    auto load_tile = [&](int kt, int stage) {
        if (tma) {
            if (tid == 0) {
                // the stage was last read through the generic proxy (LDS) before the __syncthreads that precedes this call
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                const uint32_t bar = bar_u + 8u * stage;
                mbar_expect_tx(bar, (uint32_t)(KT * (BM + BN) * 4));
                tma_load_2d(As_u + 4u * (uint32_t)(stage * KT * BM), &a.mapA, bar, (int)m0, kt * KT);
                tma_load_2d(Bs_u + 4u * (uint32_t)(stage * KT * BN), &a.mapB, bar, (int)n0, kt * KT);
            }
            return;
        }
        const uint32_t* gb = a.Bw + (int64_t)kt * KT * a.Npad + n0;
        if (CONV) {
            // The gathered A words travel through registers: a 4-byte cp.async costs one shared-memory wavefront PER
            // LANE (0.5 wavefronts per lookup for 64-column tiles, measured), a warp-wide STS of 32 consecutive words
            // one.  The loads are issued here; gather_store() writes the stage at the end of the k iteration.
            const int2* ki = a.cp.kinfo + kt * KT + tid / BM;
#pragma unroll
            for (int j = 0; j < KT * BM / NTHR; ++j) {
                const int2 q = __ldg(ki + j * (NTHR / BM));
                const int h = ch0 + (int)(short)(q.y & 0xFFFF), w = cw0 + (q.y >> 16);
                const bool ok = (unsigned)h < (unsigned)a.cp.H && (unsigned)w < (unsigned)a.cp.W;
                greg[j] = a.zero_word;
                if (ok) greg[j] = __ldg(crow + q.x);
            }
        } else {
            const uint32_t* ga = a.Aw + (int64_t)kt * KT * a.Mpad + m0;
#pragma unroll
            for (int i = tid; i < KT * BM / 4; i += NTHR) {
                const int r = i / (BM / 4), c = i % (BM / 4);
                cp_async16(As_u + 4u * (uint32_t)((stage * KT + r) * BM + c * 4), ga + (int64_t)r * a.Mpad + c * 4);
            }
        }
#pragma unroll
        for (int i = tid; i < KT * BN / 4; i += NTHR) {
            const int r = i / (BN / 4), c = i % (BN / 4);
            cp_async16(Bs_u + 4u * (uint32_t)((stage * KT + r) * BN + c * 4), gb + (int64_t)r * a.Npad + c * 4);
        }
    };

    float acc[RPW][CPL];
    int iacc[RPW][CPL];   // C8: the sums, in units of u, as integers (acc takes them over after the k loop)
#pragma unroll
    for (int i = 0; i < RPW; ++i)
#pragma unroll
        for (int j = 0; j < CPL; ++j) { acc[i][j] = 0.0f; iacc[i][j] = 0; }

    const int nk_all = a.Kpad / KT;
    constexpr bool streamk = SK;
    int kt0 = (!streamk && a.splits > 1) ? (int)((int64_t)blockIdx.z * nk_all / a.splits) : 0;
    int nk = ((!streamk && a.splits > 1) ? (int)((int64_t)(blockIdx.z + 1) * nk_all / a.splits) : nk_all) - kt0;
    float macc = 0.0f;   // split-K / stream-K: largest |partial sum| seen at a k-tile boundary
    const uint32_t zero_word = a.zero_word;
    // sampled zero statistics of the activations (predecode pass): low half zeros, high half elements
    const unsigned long long zc = *a.zero_count;
    const bool sparse = a.sparse_a >= 0 ? a.sparse_a != 0
                                        : (CPL == 4 && 10ull * (zc & 0xFFFFFFFFull) >= 4ull * (zc >> 32));
    // stream-K: this CTA's contiguous range of (tile, k-tile) units
    const int64_t sk_units = a.sk_tiles * nk_all;
    int64_t sk_u = streamk ? (int64_t)blockIdx.x * sk_units / a.sk_ctas : 0;
    const int64_t sk_end = streamk ? (int64_t)(blockIdx.x + 1) * sk_units / a.sk_ctas : 1;
    bool sk_first = true;
    uint32_t it_base = 0;   // TMA: tiles issued so far by this CTA (stage = it % ST, barrier parity = (it / ST) & 1)
  for (; sk_u < sk_end; sk_first = false) {
    if (streamk) {
        const int64_t tile = sk_u / nk_all;
        kt0 = (int)(sk_u - tile * nk_all);
        nk = (int)((nk_all - kt0 < sk_end - sk_u) ? (nk_all - kt0) : (sk_end - sk_u));
        sk_u += nk;
        m0 = (tile / a.sk_tiles_n) * BM;
        n0 = (tile % a.sk_tiles_n) * BN;
        if (CONV) conv_row_setup();
#pragma unroll
        for (int i = 0; i < RPW; ++i)
#pragma unroll
            for (int j = 0; j < CPL; ++j) { acc[i][j] = 0.0f; iacc[i][j] = 0; }
        macc = 0.0f;
        __syncthreads();   // every warp is done with the previous segment's operand stages
    } else {
        sk_u = sk_end;
    }
#pragma unroll
    for (int s = 0; s < ST - 1; ++s) {
        if (s < nk) load_tile(kt0 + s, tma ? (int)((it_base + s) % ST) : s);
        if (CONV && s < nk) gather_store(s);
        cp_async_commit();
    }
    if (C8 && kt0 == 0 && (a.esc_total[0] | *a.brare_count) != 0) {
        // the k range that starts at 0 starts from the correction sums (cells the byte table dropped)
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            const int64_t m = m0 + warp * RPW + i;
#pragma unroll
            for (int j = 0; j < CPL; ++j) {
                const int64_t n = n0 + lane * CPL + j;
                if (m < a.Mrows && n < a.N) iacc[i][j] = __ldg(a.c8_D + m * a.N + n);
            }
        }
    }
    // The k loop exists twice: dense activations use the predicated row (no branches, full ILP across
    // rows); sparse ones (post-ReLU tensors, about half zeros) use a warp-uniform branch per row.
    auto mainloop = [&](auto branchy) {
        constexpr bool BRANCHY = decltype(branchy)::value;
        for (int kt = 0; kt < nk; ++kt) {
            const uint32_t it = it_base + (uint32_t)kt;
            const int stage = tma ? (int)(it % ST) : kt % ST;
            if (tma) {
                // no CTA-wide barrier per k tile: a warp only waits for its data; warp 0 (whose lane 0 issues the
                // loads) additionally waits until every warp has released the stage it is about to refill
                mbar_wait(bar_u + 8u * stage, (it / ST) & 1u);
                if (warp == 0 && kt + ST - 1 < nk) {
                    if (it >= 1) mbar_wait(bar_u + 8u * (ST + (it - 1) % ST), ((it - 1) / ST) & 1u);
                    load_tile(kt0 + kt + ST - 1, (int)((it + ST - 1) % ST));
                }
            } else {
                cp_async_wait<ST - 2>();
                __syncthreads();
                if (kt + ST - 1 < nk) load_tile(kt0 + kt + ST - 1, (kt + ST - 1) % ST);
                cp_async_commit();
            }
            const uint32_t* Asm = As + stage * KT * BM + warp * RPW;
            const uint32_t* Bsm = Bs + stage * KT * BN + lane * CPL;
            if constexpr (!C8) {
#pragma unroll (KT >= 16 ? 4 : 2)
              for (int kk = 0; kk < KT; ++kk) {
                  uint32_t bw[CPL];
                  if (CPL == 4) {
                      const uint4 bq = *reinterpret_cast<const uint4*>(Bsm + kk * BN);
                      bw[0] = bq.x; bw[1] = bq.y; bw[CPL - 2] = bq.z; bw[CPL - 1] = bq.w;
                  } else {
                      const uint2 bq = *reinterpret_cast<const uint2*>(Bsm + kk * BN);
                      bw[0] = bq.x; bw[1] = bq.y;
                  }
                  uint32_t bo[CPL];
                  unsigned short bm[CPL];
  #pragma unroll
                  for (int j = 0; j < CPL; ++j) {
                      bo[j] = (SIDX ? bw[j] : (bw[j] & 0xFFFFu)) + lut_base;   // table byte address of the (tB_idx, mB) part
                      bm[j] = SIDX ? (unsigned short)kBf16One : (unsigned short)(bw[j] >> 16);   // bf16 sign of b
                  }
  #pragma unroll
                  for (int i4 = 0; i4 < RPW / 4; ++i4) {
                      const uint4 aq = *reinterpret_cast<const uint4*>(Asm + kk * BM + i4 * 4);
                      const uint32_t aw[4] = {aq.x, aq.y, aq.z, aq.w};
  #pragma unroll
                      for (int ii = 0; ii < 4; ++ii) {
                          if (BRANCHY) {
                              // sparse activations: a real (warp-uniform) branch, so that a zero costs neither
                              // issue slots nor LSU slots (a predicated-off LDS still takes its LSU slot)
                              if (aw[ii] != zero_word) {
  #pragma unroll
                                  for (int j = 0; j < CPL; ++j)
                                      fma_bf16(acc[i4 * 4 + ii][j], lds_u16(aw[ii] + bo[j]), bm[j]);
                              }
                          } else if (CPL == 4) mac_row4(acc[i4 * 4 + ii][0], acc[i4 * 4 + ii][1], acc[i4 * 4 + ii][CPL - 2],
                                                 acc[i4 * 4 + ii][CPL - 1], aw[ii], zero_word, bo[0], bo[1], bo[CPL - 2],
                                                 bo[CPL - 1], bm[0], bm[1], bm[CPL - 2], bm[CPL - 1]);
                          else mac_row2(acc[i4 * 4 + ii][0], acc[i4 * 4 + ii][1], aw[ii], zero_word, bo[0], bo[1],
                                                 bm[0], bm[1]);
                      }
                  }
              }
            } else {
                // byte cells: address = (A word >> 1) + (B offset >> 1), LDS.S8, IMAD by the sign of b
#pragma unroll (KT >= 16 ? 4 : 2)
                for (int kk = 0; kk < KT; ++kk) {
                    uint32_t bw[CPL], bo8[CPL];
                    int sg[CPL];
                    if (CPL == 4) {
                        const uint4 bq = *reinterpret_cast<const uint4*>(Bsm + kk * BN);
                        bw[0] = bq.x; bw[1] = bq.y; bw[CPL - 2] = bq.z; bw[CPL - 1] = bq.w;
                    } else {
                        const uint2 bq = *reinterpret_cast<const uint2*>(Bsm + kk * BN);
                        bw[0] = bq.x; bw[1] = bq.y;
                    }
#pragma unroll
                    for (int j = 0; j < CPL; ++j) {
                        bo8[j] = ((SIDX ? bw[j] : (bw[j] & 0xFFFFu)) >> 1) + lut_base;   // (tB_idx, mB) part, byte cells
                        sg[j] = SIDX ? 1 : (((int)bw[j] >> 30) | 1);   // the high half is bf16 +-1.0 (0x3F80 / 0xBF80): +1 / -1
                    }
#pragma unroll
                    for (int i4 = 0; i4 < RPW / 4; ++i4) {
                        const uint4 aq = *reinterpret_cast<const uint4*>(Asm + kk * BM + i4 * 4);
                        const uint32_t aw[4] = {aq.x, aq.y, aq.z, aq.w};
#pragma unroll
                        for (int ii = 0; ii < 4; ++ii) {
                            if (BRANCHY) {
                                if (aw[ii] != zero_word) {
                                    const uint32_t h = aw[ii] >> 1;
#pragma unroll
                                    for (int j = 0; j < CPL; ++j) iacc[i4 * 4 + ii][j] += lds_s8(h + bo8[j]) * sg[j];
                                }
                            } else if (CPL == 4) mac8_row4(iacc[i4 * 4 + ii][0], iacc[i4 * 4 + ii][1], iacc[i4 * 4 + ii][CPL - 2],
                                                   iacc[i4 * 4 + ii][CPL - 1], aw[ii], zero_word, bo8[0], bo8[1], bo8[CPL - 2],
                                                   bo8[CPL - 1], sg[0], sg[1], sg[CPL - 2], sg[CPL - 1]);
                            else mac8_row2(iacc[i4 * 4 + ii][0], iacc[i4 * 4 + ii][1], aw[ii], zero_word, bo8[0], bo8[1],
                                           sg[0], sg[1]);
                        }
                    }
                }
            }
            if (!C8 && a.splits > 1) {
#pragma unroll
                for (int i = 0; i < RPW; ++i)
#pragma unroll
                    for (int j = 0; j < CPL; ++j) macc = fmaxf(macc, fabsf(acc[i][j]));
            }
            // CONV: the tile gathered at the top of this iteration goes to its stage (last read in iteration kt - 1,
            // behind this iteration's barrier; next read two barriers from now)
            if (CONV && kt + ST - 1 < nk) gather_store((kt + ST - 1) % ST);
            if (tma) {   // this warp is done with the stage
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_u + 8u * (ST + stage)) : "memory");
            }
        }
    };
    if (sparse) mainloop(std::true_type{});
    else            mainloop(std::false_type{});
    cp_async_wait<0>();
    it_base += (uint32_t)nk;
    if constexpr (C8) {   // |sum| < 2^24 (c8_selected): the conversion is exact
#pragma unroll
        for (int i = 0; i < RPW; ++i)
#pragma unroll
            for (int j = 0; j < CPL; ++j) acc[i][j] = (float)iacc[i][j];
    }

    if (streamk && nk < nk_all) {
        // a partial k range of this tile: certificate as for split-K (see below), then the slot store
        const int t_max = a.tmax[0] + a.tmax[1] - a.t_bias;
        int lg = 0;
        while ((1 << lg) < KT) ++lg;
        const int dl = t_max + a.M + 1 + lg;
        if (!C8 && (dl > 23 || !(macc * (float)a.splits <= 16777216.0f - (float)(1 << (dl > 0 ? dl : 0))))) *a.inexact = 1;
        float* P = a.partial + ((int64_t)blockIdx.x * 2 + (sk_first ? 0 : 1)) * (int64_t)(BM * BN);
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            float* prow = P + (warp * RPW + i) * BN + lane * CPL;
            if (CPL == 4) *reinterpret_cast<float4*>(prow) = make_float4(acc[i][0], acc[i][1], acc[i][CPL - 2], acc[i][CPL - 1]);
            else          *reinterpret_cast<float2*>(prow) = make_float2(acc[i][0], acc[i][1]);
        }
        continue;
    }
    if (!streamk && a.splits > 1) {
        // Exactness certificate.  Table entries are integers (units of u) of magnitude <= 2^(t + M + 1), so one
        // k tile moves a running sum by at most D = KT * 2^(t_max + M + 1).  If every slice's running sum stayed
        // within (2^24 - D) / splits at all of its k-tile boundaries, then every intermediate sum of every slice,
        // of the ordered combination of the slices AND of the unsplit k-ascending order is an integer of
        // magnitude <= 2^24: all of them are exact in fp32, hence the result is bit-identical to the unsplit
        // kernel's.
        const int t_max = a.tmax[0] + a.tmax[1] - a.t_bias;
        int lg = 0;
        while ((1 << lg) < KT) ++lg;
        const int dl = t_max + a.M + 1 + lg;   // log2 of D, the most one k tile can move a sum
        // (C8: every sum of table cells is an integer below 2^24 under c8_selected(): exact in any order)
        if (!C8 && (dl > 23 || !(macc * (float)a.splits <= 16777216.0f - (float)(1 << (dl > 0 ? dl : 0))))) *a.inexact = 1;
        float* P = a.partial + (int64_t)blockIdx.z * a.Mrows * a.N;
        const int64_t n = n0 + lane * CPL;
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            const int64_t m = m0 + warp * RPW + i;
            if (m >= a.Mrows) break;
#pragma unroll
            for (int j = 0; j < CPL; ++j)
                if (n + j < a.N) P[m * a.N + n + j] = acc[i][j];
        }
        return;
    }
    // epilogue: C = acc * u (exact power-of-two scaling); quantize_after_mult_and_add: C = Q(acc * u)
    const float unit = pow2f_i(1 - a.bias.R() - a.M);
    if (a.qamaa) {
        QamaaOp q = a.qop;
        q.begin();
#pragma unroll
        for (int i = 0; i < RPW; ++i)
#pragma unroll
            for (int j = 0; j < CPL; ++j) acc[i][j] = q.quant(acc[i][j] * unit) / unit;   // stored below as acc * unit
    }
    const int64_t n = n0 + lane * CPL;
    if (a.ep.on) {
        // fused epilogue (epilogue.cuh): the lane owns CPL fixed columns = output channels, so their bias / BN
        // constants live in registers; the final values are computed here
        Epilogue::Col ec[CPL];
#pragma unroll
        for (int j = 0; j < CPL; ++j) ec[j] = a.ep.column(n + j < a.N ? n + j : a.N - 1);
#pragma unroll
        for (int i = 0; i < RPW; ++i)
#pragma unroll
            for (int j = 0; j < CPL; ++j) acc[i][j] = a.ep.apply(acc[i][j] * unit, ec[j]);
        if (a.ep.HW > 0) {
            // NCHW store.  A warp holds RPW consecutive rows (pixels) of 32 * CPL channels: transposed through a
            // warp-private patch of the (now idle) operand ring, one column of every lane per pass, so that a
            // store instruction writes runs of RPW consecutive pixels (RPW * 4 contiguous bytes per channel).
            static_assert(RPW == 16 || RPW == 32, "the transposing store assumes 16 or 32 rows per warp");
            static_assert((size_t)ST * KT * (BM + BN) * 4 >= (size_t)WARPS * 32 * (RPW + 1) * 4, "staging patch fits the operand ring");
            __syncthreads();   // every warp is done reading operand stages (no cp.async is in flight: wait<0> above)
            float* patch = reinterpret_cast<float*>(As) + warp * (32 * (RPW + 1));
            const int row = lane % RPW, cgrp = lane / RPW;           // this lane's row in the store phase
            constexpr int CPI = 32 / RPW;                            // channels per store instruction
            const int64_t m = m0 + warp * RPW + row;
            const int64_t bimg = m / a.ep.HW, pix = m - bimg * a.ep.HW;
            float* obase = a.C + (bimg * a.ep.Ctot + a.ep.c0) * a.ep.HW + pix;
#pragma unroll
            for (int j = 0; j < CPL; ++j) {
#pragma unroll
                for (int i = 0; i < RPW; ++i) patch[lane * (RPW + 1) + i] = acc[i][j];
                __syncwarp();
                if (m < a.Mrows) {
#pragma unroll 4
                    for (int c = cgrp; c < 32; c += CPI) {
                        const int64_t nn = n0 + c * CPL + j;
                        if (nn < a.N) obase[nn * a.ep.HW] = patch[c * (RPW + 1) + row];
                    }
                }
                __syncwarp();
            }
            if (streamk) __syncthreads();   // the next segment's prologue overwrites the ring
            continue;
        }
        // row-major: the plain store below, with the scaling already applied
        const bool vec = ((a.ldc & (CPL - 1)) == 0) && ((reinterpret_cast<uintptr_t>(a.C) & (4 * CPL - 1)) == 0) &&
                         (n + CPL - 1 < a.N);
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            const int64_t m = m0 + warp * RPW + i;
            if (m >= a.Mrows) break;
            float* crow = a.C + m * a.ldc;
            if (vec && CPL == 4) {
                *reinterpret_cast<float4*>(crow + n) = make_float4(acc[i][0], acc[i][1], acc[i][CPL - 2], acc[i][CPL - 1]);
            } else if (vec) {
                *reinterpret_cast<float2*>(crow + n) = make_float2(acc[i][0], acc[i][1]);
            } else {
#pragma unroll
                for (int j = 0; j < CPL; ++j)
                    if (n + j < a.N) crow[n + j] = acc[i][j];
            }
        }
        continue;
    }
    const bool vec = ((a.ldc & (CPL - 1)) == 0) && ((reinterpret_cast<uintptr_t>(a.C) & (4 * CPL - 1)) == 0) &&
                     (n + CPL - 1 < a.N);
#pragma unroll
    for (int i = 0; i < RPW; ++i) {
        const int64_t m = m0 + warp * RPW + i;
        if (m >= a.Mrows) break;
        float* crow = a.C + m * a.ldc;
        if (vec && CPL == 4) {
            *reinterpret_cast<float4*>(crow + n) =
                make_float4(acc[i][0] * unit, acc[i][1] * unit, acc[i][CPL - 2] * unit, acc[i][CPL - 1] * unit);
        } else if (vec) {
            *reinterpret_cast<float2*>(crow + n) = make_float2(acc[i][0] * unit, acc[i][1] * unit);
        } else {
#pragma unroll
            for (int j = 0; j < CPL; ++j)
                if (n + j < a.N) crow[n + j] = acc[i][j] * unit;
        }
    }
  }   // stream-K segment loop (one pass otherwise)
}

// byte-cell instantiations live in approx_gemm_c8.cu: kernel and stream-K kernel for a tile configuration (false: none)
bool lut_c8_kernels(int warps, int rpw, int cpl, int kt, int sidx, void (**kernel)(LutGemmArgs), void (**sk_kernel)(LutGemmArgs));
int lut_c8_set_attributes();   // opt-in shared-memory size of every byte-cell kernel, for the current device

}  // namespace b200
