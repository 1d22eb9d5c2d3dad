// fq_gemm_tc.cu -- K3: the NON-approx matmul of fake-quantized operands on the 5th-generation tensor
// cores.
//
// Replaces `x @ y[:, i:i+1]` per column (approx/approx_calculation.py:968-969, :796-797, :996) and
// `self.multiply` = torch.matmul (:704-705, conv without BN).
//
// Why tensor cores are legal here and only here: with approx_flag == False there is no per-product
// bit manipulation and no re-quantisation between the multiply and the add -- the layer is a plain
// contraction of fake-quantized operands.  Those operands carry <= 6 significant bits (mant_width <= 5
// plus the hidden one) and an fp32-range exponent, so they convert to bf16 EXACTLY; bf16 x bf16
// products are exact in fp32; only the fp32 accumulation order differs from the reference's
// (cuBLAS/MKL sgemm, itself unspecified).  Stated tolerance: |C - ref| <= 1e-6 * sum_k |a_k b_k|.
// Hardware FP8 (kind::f8f6f4) is NOT used: e4m3/e5m2 reserve exponent codes for NaN/Inf while the
// reference's grid (fp8_quantizer.py:13-41) uses all of them and moves the bias per channel.
//
// Pipeline:  convert pass (fp32 -> bf16 planes, K padded to 64, raises a flag if ANY operand is not
// exactly representable)  ->  fq_gemm_tc_kernel: TMA (128B swizzle) -> 4-stage shared-memory ring ->
// tcgen05.mma kind::f16 issued by one thread, fp32 accumulator in TMEM -> tcgen05.ld epilogue
// (+bias) -> fp32 stores.  If the flag is up (operands not fake-quantized), the tensor-core kernel
// exits at once and an fp32 CUDA-core kernel computes the product instead: still no CPU path.
#include <cuda.h>
#include <cstring>
#include <cuda_bf16.h>

#include <algorithm>
#include <mutex>

#include "common.cuh"
#include "gemm_impl.cuh"
#include "pair_gemm.cuh"

namespace b200 {

constexpr int TC_BM = 128;      // UMMA M (cta_group::1): TMEM lane i <-> tile row i
constexpr int TC_BK = 64;       // one 128-byte swizzle row of bf16
constexpr int TC_UK = 16;       // K per tcgen05.mma kind::f16
constexpr int TC_THREADS = 256; // warp 0: TMA, warp 1: MMA, warp 2: TMEM alloc, warps 4-7: epilogue

// ---- PTX wrappers ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
        "@P1 bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(s32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(s32(dst)), "l"(map), "r"(s32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 operands, fp32 accumulate
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = TMEM lane = tile row)
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// A fake-quantizer evaluated inside this file's kernels (opt-in fusions): on the left operand while it is converted
// to bf16 (the hijacker's `x = self.activation_quantizer(x)`, quantization/hijacker.py:82-88) and on the result in the
// epilogue (`res = self.res_quantizer(res)`, :93-94).  Same element function as the stand-alone kernel
// (quantize_to_fp8_ste_MM, quantization/quantizers/fp8_quantizer.py:119-154), with the exponent bias that kernel
// derived from the same per-tensor maxval: bit-identical to the separate pass.
struct FqParams {
    const float* maxval;   // device [1] or null: no quantizer
    const float* bias;     // device [1]
    float Mf;
    int sign;
    __device__ __forceinline__ float apply(float v, float mx, float fb, bool fast) const
    {
        return fast ? fp8_ste_quant_fast(v, mx, (int)fb, (int)Mf, sign) : fp8_ste_quant(v, mx, fb, Mf, sign);
    }
};

// Epilogue store of one 32 x 32 fp32 chunk through shared memory and TMA: thread = row, v = its 32 consecutive
// columns.  The chunk is written in the 128-byte-swizzle layout the C tensor map expects (16-byte chunk q of row r at
// chunk q ^ (r & 7): conflict-free 128-bit shared stores), then ONE lane issues cp.async.bulk.tensor (global <- shared):
// full 128-byte lines instead of 16 bytes per row per store instruction, ragged edges clipped by the tensor map.
// `buf`: this warp's 4 KB staging buffer for this chunk parity (1024-byte aligned).
constexpr int TC_EPI_BYTES = 4 * 2 * 4096;     // 4 epilogue warps x double buffer x (32 rows x 128 B)
template <int PER>
__device__ __forceinline__ void epi_store_chunk_tma(const CUtensorMap* mapC, uint8_t* buf, const uint32_t (&v)[32], int lane,
                                                    bool has_bias, const float (&breg)[PER], int c, int64_t n, int64_t N, int row0,
                                                    const FqParams& oq, float omx, float ofb, bool ofast)
{
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store that last read `buf` is done with it
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        float4 o = make_float4(__uint_as_float(v[4 * q]), __uint_as_float(v[4 * q + 1]), __uint_as_float(v[4 * q + 2]),
                               __uint_as_float(v[4 * q + 3]));
        if (has_bias) {
            // lane L holds the tile's bias values L * PER .. L * PER + PER - 1 (0 beyond N), fetched before the wait for
            // the accumulator; column 4q + i of chunk c is value (c * 32 + 4q + i): static register, shuffled lane
            const int col = c * 32 + 4 * q;
            o.x += __shfl_sync(0xFFFFFFFFu, breg[(4 * q) % PER], col / PER);
            o.y += __shfl_sync(0xFFFFFFFFu, breg[(4 * q + 1) % PER], (col + 1) / PER);
            o.z += __shfl_sync(0xFFFFFFFFu, breg[(4 * q + 2) % PER], (col + 2) / PER);
            o.w += __shfl_sync(0xFFFFFFFFu, breg[(4 * q + 3) % PER], (col + 3) / PER);
        }
        if (oq.maxval) {
            o.x = oq.apply(o.x, omx, ofb, ofast); o.y = oq.apply(o.y, omx, ofb, ofast);
            o.z = oq.apply(o.z, omx, ofb, ofast); o.w = oq.apply(o.w, omx, ofb, ofast);
        }
        *reinterpret_cast<float4*>(buf + lane * 128 + ((q ^ (lane & 7)) << 4)) = o;
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (lane == 0) {
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                     ::"l"(mapC), "r"(s32(buf)), "r"((int)n), "r"(row0) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
}

// Shared-memory matrix descriptor of a K-major tile stored as TMA's 128-byte swizzle writes it:
// row r at byte r*128, 16-byte chunks XOR-ed with (r & 7); 8-row groups are 1024 B apart (SBO).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr)
{
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);          // start address, 16-byte units
    d |= (uint64_t)0 << 16;                               // leading byte offset: unused for swizzled K-major
    d |= (uint64_t)((1024u >> 4) & 0x3FFFu) << 32;        // stride byte offset: 8 rows * 128 B
    d |= (uint64_t)1 << 46;                               // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                               // layout type: SWIZZLE_128B
    return d;
}
// Instruction descriptor, kind::f16: D fp32, A/B bf16, both K-major, M x N tile.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N)
{
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- the tensor-core kernel --------------------------------------------------------------------
template <int BN, int ST>
struct TcSmem {
    static constexpr int A_BYTES = TC_BM * TC_BK * 2;        // 16 KB
    static constexpr int B_BYTES = BN * TC_BK * 2;
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int BARS = ST * STAGE;                  // full[ST], empty[ST], tmem_full[2], tmem_empty[2], TMEM base
    static constexpr int EPI = BARS + 1024;                  // epilogue staging (TMA stores), 1024-byte aligned
    static constexpr int TOTAL = 1024 + BARS + 1024 + TC_EPI_BYTES;   // alignment slack, stages, barriers, staging
};

__device__ __forceinline__ void mbar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}

// Persistent: one CTA per SM walks the output tiles (m fastest, so that concurrently running CTAs share
// the B tile in L2).  Three pipelines run concurrently: TMA -> shared-memory ring (full/empty), MMA ->
// one of TWO accumulator stages in tensor memory (tmem_full/tmem_empty), epilogue warps draining the
// other stage -- the epilogue of tile i overlaps the main loop of tile i+1.
template <int BN, int ST>
__global__ void __launch_bounds__(TC_THREADS, 1)
fq_gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                  const __grid_constant__ CUtensorMap mapC, int tma_store,
                  float* __restrict__ C, int64_t ldc, int64_t Mrows, int64_t N, int nkb, int m_blocks, int n_tiles,
                  const float* __restrict__ bias, const int* __restrict__ inexact, FqParams oq)
{
    if (*inexact != 0) return;   // an operand does not fit bf16 exactly: the fp32 kernel takes over
    using L = TcSmem<BN, ST>;
    constexpr int TMEM_COLS = 2 * BN <= 256 ? 256 : 512;   // two accumulator stages; allocations are powers of two
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::BARS);
    uint64_t* empty = full + ST;
    uint64_t* tmem_full = empty + ST;      // [2]
    uint64_t* tmem_empty = tmem_full + 2;  // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < ST; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {   // one warp allocates two accumulator stages of BN fp32 columns each
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "r"((uint32_t)TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ===== TMA producer =====
        int it = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int m0 = (tile % m_blocks) * TC_BM, n0 = (tile / m_blocks) * BN;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % ST;
                mbar_wait(empty + s, ((it / ST) & 1) ^ 1);
                mbar_expect_tx(full + s, L::STAGE);
                tma_load_2d(smem + s * L::STAGE, &mapA, full + s, kb * TC_BK, m0);
                tma_load_2d(smem + s * L::STAGE + L::A_BYTES, &mapB, full + s, kb * TC_BK, n0);
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ===== MMA issuer: one thread drives the tensor core for the whole CTA =====
        constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, BN);
        int it = 0, ti = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ti) {
            const int as = ti & 1;
            mbar_wait(tmem_empty + as, ((ti >> 1) & 1) ^ 1);     // the epilogue has drained this accumulator stage
            tc_fence_after();
            const uint32_t d_tmem = tmem + (uint32_t)(as * BN);
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % ST;
                mbar_wait(full + s, (it / ST) & 1);
                tc_fence_after();
                const uint64_t adesc = umma_desc_sw128(s32(smem + s * L::STAGE));
                const uint64_t bdesc = umma_desc_sw128(s32(smem + s * L::STAGE + L::A_BYTES));
#pragma unroll
                for (int k = 0; k < TC_BK / TC_UK; ++k)   // +32 B (= 2 sixteen-byte units) per K step inside the swizzle row
                    tc_mma_bf16(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
                tc_commit(empty + s);      // frees the stage when these MMAs have read it
            }
            tc_commit(tmem_full + as);     // accumulator complete
        }
    } else if (warp >= 4) {
        // ===== epilogue: TMEM -> registers -> (+bias) -> global fp32 =====
        const int w = warp - 4;                       // == warp % 4: the TMEM lane quarter this warp may read
        const bool vec_ok = ((ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0) &&
                            ((reinterpret_cast<uintptr_t>(bias) & 15) == 0);   // bias is read as float4 beside C
        const float omx = oq.maxval ? __ldg(oq.maxval) : 0.0f, ofb = oq.maxval ? __ldg(oq.bias) : 0.0f;   // result quantizer
        const bool ofast = oq.maxval && fp8_ste_fast_ok(ofb);
        int ti = 0;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++ti) {
            const int m0 = (tile % m_blocks) * TC_BM, n0 = (tile / m_blocks) * BN;
            const int as = ti & 1;
            // The tile's bias goes to registers (lane L: the tile's columns L * PER ...) BEFORE the wait for the
            // accumulator: a global load per 32-column chunk sat on the epilogue's critical path (+20 us on
            // 4096 x 768 x 3072 with a bias); the chunks hand the values round by shuffle.
            // (PER values per lane over a 32 * PER-wide window from n0; PER is 4 or 8 so that 32 % PER == 0 and the
            // register a chunk column needs does not depend on the chunk -- a 192-wide tile uses the 256-wide window)
            constexpr int PER = BN <= 128 ? 4 : 8;
            float breg[PER];
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                const int64_t nn = (int64_t)n0 + lane * PER + i;
                breg[i] = (bias && tma_store && nn < N) ? __ldg(bias + nn) : 0.0f;
            }
            mbar_wait(tmem_full + as, (ti >> 1) & 1);
            tc_fence_after();
            const int64_t row = (int64_t)m0 + w * 32 + lane;
            float* crow = C + row * ldc;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tc_ld32(tmem + ((uint32_t)(w * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                const int64_t n = (int64_t)n0 + c * 32;
                if (tma_store) {
                    if (n < N && (int64_t)m0 + w * 32 < Mrows)
                        epi_store_chunk_tma<PER>(&mapC, smem + L::EPI + (w * 2 + (c & 1)) * 4096, v, lane, bias != nullptr, breg, c, n, N,
                                            m0 + w * 32, oq, omx, ofb, ofast);
                } else if (row < Mrows && n < N) {
                    if (vec_ok && n + 32 <= N) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float4 o = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                   __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                            if (bias) {
                                const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + n + j));
                                o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
                            }
                            if (oq.maxval) {
                                o.x = oq.apply(o.x, omx, ofb, ofast); o.y = oq.apply(o.y, omx, ofb, ofast);
                                o.z = oq.apply(o.z, omx, ofb, ofast); o.w = oq.apply(o.w, omx, ofb, ofast);
                            }
                            *reinterpret_cast<float4*>(crow + n + j) = o;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n + j < N) {
                                float o = __uint_as_float(v[j]);
                                if (bias) o += __ldg(bias + n + j);
                                crow[n + j] = oq.maxval ? oq.apply(o, omx, ofb, ofast) : o;
                            }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + as);    // this warp's quarter of the stage is drained
        }
        if (tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores complete before exit
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)TMEM_COLS) : "memory");
}

// ---- CTA-pair variant (cta_group::2) ------------------------------------------------------------------------------
// Two CTAs of a cluster (the two SMs of a TPC) compute ONE 256 x 256 output tile: each loads its own 128 rows of A and
// its own 128 columns (rows of the [N][K] plane) of B; the LEADER CTA's single MMA thread issues
// tcgen05.mma.cta_group::2 (M = 256, N = 256), which reads A from each CTA's shared memory for that CTA's half of the
// accumulator rows and B from BOTH CTAs' shared memory.  Per SM and k block that is 16 KB of A + 16 KB of B instead of
// 16 + 32: the shared-memory operand feed -- what holds the single-CTA kernel at ~55 % tensor-pipe activity -- is
// halved, and six pipeline stages fit where four did.
//   full[s]   (leader): 1 arrival (leader's producer: expect_tx of BOTH CTAs' bytes) + the TMA bytes of both CTAs
//   empty[s]  (each CTA): tcgen05.commit multicast to both CTAs when the MMAs that read stage s have finished
//   tmem_full[a]  (each CTA): commit multicast: accumulator stage a is complete
//   tmem_empty[a] (leader): 8 arrivals = 4 epilogue warps of each CTA (the peer's arrive remotely)
constexpr int TC2_BN = 256, TC2_ST = 6, TC2_HALF_N = 128;
struct Tc2Smem {
    static constexpr int A_BYTES = TC_BM * TC_BK * 2;          // 16 KB: this CTA's 128 rows
    static constexpr int B_BYTES = TC2_HALF_N * TC_BK * 2;     // 16 KB: this CTA's 128 columns of B
    static constexpr int STAGE = A_BYTES + B_BYTES;
    static constexpr int BARS = TC2_ST * STAGE;
    static constexpr int EPI = BARS + 1024;
    static constexpr int TOTAL = 1024 + BARS + 1024 + TC_EPI_BYTES;
};

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// both CTAs execute this; the byte count lands on the LEADER's barrier (same offset, CTA rank 0)
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* map, uint32_t leader_bar, int c0, int c1)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(s32(dst)), "l"(map), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ uint32_t map_to_rank0(uint32_t local_addr)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(r) : "r"(local_addr));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr)
{
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(s32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_pair(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_THREADS, 1)
fq_gemm_tc2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                   const __grid_constant__ CUtensorMap mapC, int tma_store,
                   float* __restrict__ C, int64_t ldc, int64_t Mrows, int64_t N, int nkb, int m_blocks, int n_tiles,
                   const float* __restrict__ bias, const int* __restrict__ inexact, FqParams oq)
{
    if (*inexact != 0) return;   // uniform over the grid: both CTAs of a pair leave together
    using L = Tc2Smem;
    constexpr int ST = TC2_ST, BN = TC2_BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (s32(smem_raw) & 1023u)) & 1023u);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::BARS);
    uint64_t* empty = full + ST;
    uint64_t* tmem_full = empty + ST;      // [2]
    uint64_t* tmem_empty = tmem_full + 2;  // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < ST; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {   // the same logical warp in both CTAs: 2 accumulator stages x 256 columns = all of tensor memory
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "r"((uint32_t)(2 * BN)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // the peer's barriers exist before anything arrives on them
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ===== TMA producer (both CTAs): own rows of A, own columns of B; bytes are counted on the leader's barrier =====
        int it = 0;
        for (int tile = pair; tile < n_tiles; tile += n_pairs) {
            const int m0 = (tile % m_blocks) * (2 * TC_BM) + (int)rank * TC_BM;
            const int n0 = (tile / m_blocks) * BN + (int)rank * TC2_HALF_N;
            for (int kb = 0; kb < nkb; ++kb, ++it) {
                const int s = it % ST;
                mbar_wait(empty + s, ((it / ST) & 1) ^ 1);
                const uint32_t lbar = map_to_rank0(s32(full + s));
                if (leader) mbar_expect_tx(full + s, 2 * L::STAGE);
                tma_load_2d_pair(smem + s * L::STAGE, &mapA, lbar, kb * TC_BK, m0);
                tma_load_2d_pair(smem + s * L::STAGE + L::A_BYTES, &mapB, lbar, kb * TC_BK, n0);
            }
        }
    } else if (warp == 1 && lane == 0) {
        if (leader) {
            // ===== MMA issuer: one thread of the leader CTA drives both SMs' tensor cores =====
            constexpr uint32_t idesc = umma_idesc_bf16(2 * TC_BM, BN);
            int it = 0, ti = 0;
            for (int tile = pair; tile < n_tiles; tile += n_pairs, ++ti) {
                const int as = ti & 1;
                mbar_wait(tmem_empty + as, ((ti >> 1) & 1) ^ 1);     // both CTAs' epilogues have drained this stage
                tc_fence_after();
                const uint32_t d_tmem = tmem + (uint32_t)(as * BN);
                for (int kb = 0; kb < nkb; ++kb, ++it) {
                    const int s = it % ST;
                    mbar_wait(full + s, (it / ST) & 1);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(s32(smem + s * L::STAGE));
                    const uint64_t bdesc = umma_desc_sw128(s32(smem + s * L::STAGE + L::A_BYTES));
#pragma unroll
                    for (int k = 0; k < TC_BK / TC_UK; ++k)
                        tc_mma_bf16_pair(d_tmem, adesc + 2u * k, bdesc + 2u * k, idesc, (kb | k) != 0);
                    tc_commit_pair(empty + s);     // frees stage s in BOTH CTAs
                }
                tc_commit_pair(tmem_full + as);    // accumulator complete, in both CTAs
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue (both CTAs): own 128 rows x 256 columns, TMEM -> registers -> (+bias) -> global fp32 =====
        const int w = warp - 4;
        const bool vec_ok = ((ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15) == 0) &&
                            ((reinterpret_cast<uintptr_t>(bias) & 15) == 0);   // bias is read as float4 beside C
        const float omx = oq.maxval ? __ldg(oq.maxval) : 0.0f, ofb = oq.maxval ? __ldg(oq.bias) : 0.0f;   // result quantizer
        const bool ofast = oq.maxval && fp8_ste_fast_ok(ofb);
        int ti = 0;
        for (int tile = pair; tile < n_tiles; tile += n_pairs, ++ti) {
            const int m0 = (tile % m_blocks) * (2 * TC_BM) + (int)rank * TC_BM, n0 = (tile / m_blocks) * BN;
            const int as = ti & 1;
            // The tile's bias goes to registers (lane L: the tile's columns L * PER ...) BEFORE the wait for the
            // accumulator: a global load per 32-column chunk sat on the epilogue's critical path (+20 us on
            // 4096 x 768 x 3072 with a bias); the chunks hand the values round by shuffle.
            // (PER values per lane over a 32 * PER-wide window from n0; PER is 4 or 8 so that 32 % PER == 0 and the
            // register a chunk column needs does not depend on the chunk -- a 192-wide tile uses the 256-wide window)
            constexpr int PER = BN <= 128 ? 4 : 8;
            float breg[PER];
#pragma unroll
            for (int i = 0; i < PER; ++i) {
                const int64_t nn = (int64_t)n0 + lane * PER + i;
                breg[i] = (bias && tma_store && nn < N) ? __ldg(bias + nn) : 0.0f;
            }
            mbar_wait(tmem_full + as, (ti >> 1) & 1);
            tc_fence_after();
            const int64_t row = (int64_t)m0 + w * 32 + lane;
            float* crow = C + row * ldc;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                tc_ld32(tmem + ((uint32_t)(w * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                const int64_t n = (int64_t)n0 + c * 32;
                if (tma_store) {
                    if (n < N && (int64_t)m0 + w * 32 < Mrows)
                        epi_store_chunk_tma<PER>(&mapC, smem + L::EPI + (w * 2 + (c & 1)) * 4096, v, lane, bias != nullptr, breg, c, n, N,
                                            m0 + w * 32, oq, omx, ofb, ofast);
                } else if (row < Mrows && n < N) {
                    if (vec_ok && n + 32 <= N) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float4 o = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]),
                                                   __uint_as_float(v[j + 2]), __uint_as_float(v[j + 3]));
                            if (bias) {
                                const float4 bv = __ldg(reinterpret_cast<const float4*>(bias + n + j));
                                o.x += bv.x; o.y += bv.y; o.z += bv.z; o.w += bv.w;
                            }
                            if (oq.maxval) {
                                o.x = oq.apply(o.x, omx, ofb, ofast); o.y = oq.apply(o.y, omx, ofb, ofast);
                                o.z = oq.apply(o.z, omx, ofb, ofast); o.w = oq.apply(o.w, omx, ofb, ofast);
                            }
                            *reinterpret_cast<float4*>(crow + n + j) = o;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n + j < N) {
                                float o = __uint_as_float(v[j]);
                                if (bias) o += __ldg(bias + n + j);
                                crow[n + j] = oq.maxval ? oq.apply(o, omx, ofb, ofast) : o;
                            }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(map_to_rank0(s32(tmem_empty + as)));   // counted on the leader
        }
        if (tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // nobody leaves while the pair's barriers / shared memory may still be signalled
    if (warp == 2)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)(2 * BN)) : "memory");
}

// ---- convert pass: fp32 operand -> bf16 plane [rows][Kp], k contiguous, zero padded to Kp ---------
// element (row, k) at src[row * s_row + k * s_k].  Raises *inexact if a value loses bits in bf16.
template <bool SRC_K_CONTIG>
__global__ void __launch_bounds__(256) to_bf16_kernel(const float* __restrict__ src, int64_t s_row, int64_t s_k,
                                                      __nv_bfloat16* __restrict__ dst, int64_t rows, int64_t K,
                                                      int64_t Kp, int* __restrict__ inexact, FqParams iq)
{
    int bad = 0;
    const float imx = iq.maxval ? __ldg(iq.maxval) : 0.0f, ifb = iq.maxval ? __ldg(iq.bias) : 0.0f;   // fused input quantizer
    const bool ifast = iq.maxval && fp8_ste_fast_ok(ifb);
    if (SRC_K_CONTIG) {
        // one thread: 8 consecutive k of one row (two float4 loads when the source allows it, one 16-byte store)
        const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        const int64_t per_row = Kp / 8;
        if (q < rows * per_row) {
            const int64_t r = q / per_row, k = (q % per_row) * 8;
            const float* p = src + r * s_row + k;
            float v[8] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
            if (k + 7 < K && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
                const float4 t0 = __ldg(reinterpret_cast<const float4*>(p));
                const float4 t1 = __ldg(reinterpret_cast<const float4*>(p) + 1);
                v[0] = t0.x; v[1] = t0.y; v[2] = t0.z; v[3] = t0.w;
                v[4] = t1.x; v[5] = t1.y; v[6] = t1.z; v[7] = t1.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (k + j < K) v[j] = __ldg(p + j);
            }
            uint32_t h[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (iq.maxval) v[j] = iq.apply(v[j], imx, ifb, ifast);
                const uint32_t b = __float_as_uint(v[j]);
                bad |= (b & 0xFFFFu) != 0;
                h[j] = b >> 16;
            }
            *reinterpret_cast<uint4*>(dst + r * Kp + k) =
                make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
        }
    } else {
        // source contiguous along rows (a [K, N] row-major right operand): 32 x 32 transpose tile
        __shared__ uint16_t tile[32][33];
        const int64_t r0 = (int64_t)blockIdx.x * 32, k0 = (int64_t)blockIdx.y * 32;
        const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t k = k0 + ty + i * 8, r = r0 + tx;
            uint32_t b = 0;
            if (r < rows && k < K) {
                float v = __ldg(src + r * s_row + k * s_k);
                if (iq.maxval) v = iq.apply(v, imx, ifb, ifast);
                b = __float_as_uint(v);
            }
            bad |= (b & 0xFFFFu) != 0;
            tile[ty + i * 8][tx] = (uint16_t)(b >> 16);   // [k][r]
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t r = r0 + ty + i * 8, k = k0 + tx;
            if (r < rows && k < Kp) reinterpret_cast<uint16_t*>(dst)[r * Kp + k] = tile[tx][ty + i * 8];
        }
    }
    if (bad) *inexact = 1;
}

// fp32 CUDA-core product for operands that are not on a bf16-exact grid (and for K == 0)
struct PlainOp {
    struct Col { float bias; };
    const float* bias;
    FqParams oq;           // result quantizer (opt-in), or maxval == null
    float omx, ofb;
    bool ofast;
    __device__ __forceinline__ void begin()
    {
        omx = oq.maxval ? __ldg(oq.maxval) : 0.0f;
        ofb = oq.maxval ? __ldg(oq.bias) : 0.0f;
        ofast = oq.maxval && fp8_ste_fast_ok(ofb);
    }
    __device__ __forceinline__ Col column(int64_t n) const { return Col{bias ? __ldg(bias + n) : 0.0f}; }
    __device__ __forceinline__ float pair(float a, float b, const Col&) const { return a * b; }
    __device__ __forceinline__ float finish(float acc, const Col& c) const
    {
        const float o = bias ? acc + c.bias : acc;
        return oq.maxval ? oq.apply(o, omx, ofb, ofast) : o;
    }
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int get_encode_fn(EncodeTiledFn* out)
{
    static std::mutex mu;
    static EncodeTiledFn fn = nullptr;
    std::lock_guard<std::mutex> lk(mu);
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        B200_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
        if (!p || q != cudaDriverEntryPointSuccess) { set_error("cuTensorMapEncodeTiled is not available in this driver"); return B200_ERR_CUDA; }
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    *out = fn;
    return B200_OK;
}

// bf16 plane [rows][Kp] -> 2-D tensor map, box {64 (k), box_rows}, 128-byte swizzle, OOB rows read as zero
static int make_map(EncodeTiledFn enc, CUtensorMap* map, const void* base, int64_t rows, int64_t Kp, int box_rows)
{
    const cuuint64_t gdim[2] = {(cuuint64_t)Kp, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)Kp * 2};
    const cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (code " + std::to_string((int)r) + ")"); return B200_ERR_CUDA; }
    return B200_OK;
}

// C [Mrows][ldc] fp32 as a 2-D tensor {N, Mrows}, box 32 columns x 32 rows, 128-byte swizzle (the epilogue's TMA stores)
static int make_map_c(EncodeTiledFn enc, CUtensorMap* map, const float* C, int64_t ldc, int64_t Mrows, int64_t N)
{
    const cuuint64_t gdim[2] = {(cuuint64_t)N, (cuuint64_t)Mrows};
    const cuuint64_t gstride[1] = {(cuuint64_t)ldc * 4};
    const cuuint32_t box[2] = {32, 32};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(C), gdim, gstride, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled(C) failed (code " + std::to_string((int)r) + ")"); return B200_ERR_CUDA; }
    return B200_OK;
}

constexpr size_t kTcFlagBytes = 256;

}  // namespace b200

using namespace b200;

extern "C" size_t b200_fq_gemm_tc_workspace_bytes(int64_t Mrows, int64_t K, int64_t N)
{
    if (Mrows < 0 || K < 0 || N < 0) return 0;
    const int64_t Kp = round_up(K > 0 ? K : 1, TC_BK);
    return kTcFlagBytes + (size_t)round_up((Mrows + N) * Kp * 2, 256) + 256;
}

// The right operand (the layer's fake-quantized weights) is the same tensor call after call once the ranges are
// fixed: b200_fq_gemm_tc_prepare_b converts it ONCE into the bf16 plane the kernel's tensor map reads
// ([256-byte header: int32 "not exact in bf16" flag][N][Kp] bf16), b200_fq_gemm_tc_pre then skips that convert pass.
extern "C" size_t b200_fq_gemm_tc_bplane_bytes(int64_t K, int64_t N)
{
    if (K < 0 || N < 0) return 0;
    const int64_t Kp = round_up(K > 0 ? K : 1, TC_BK);
    return kTcFlagBytes + (size_t)round_up(N * Kp * 2, 256);
}

static int convert_b(const float* B, int64_t sb_k, int64_t sb_n, int64_t K, int64_t N, int64_t Kp, __nv_bfloat16* Bh, int* flag,
                     cudaStream_t st)
{
    if (sb_k == 1) {
        const int64_t n = N * (Kp / 8);
        to_bf16_kernel<true><<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(B, sb_n, 1, Bh, N, K, Kp, flag, FqParams{});
    } else {
        to_bf16_kernel<false><<<dim3((unsigned)ceil_div(N, 32), (unsigned)(Kp / 32)), 256, 0, st>>>(B, sb_n, sb_k, Bh, N, K, Kp, flag, FqParams{});
    }
    B200_LAUNCH_CHECK("to_bf16_kernel<B>");
    return B200_OK;
}

extern "C" int b200_fq_gemm_tc_prepare_b(const float* B, int64_t sb_k, int64_t sb_n, int64_t K, int64_t N, void* bplane,
                                         size_t bplane_bytes, void* stream)
{
    if (int rc = check_device()) return rc;
    if (K < 1 || N < 1 || !B || !bplane) { set_error("b200_fq_gemm_tc_prepare_b: bad argument"); return B200_ERR_ARGUMENT; }
    if (bplane_bytes < b200_fq_gemm_tc_bplane_bytes(K, N) || ((uintptr_t)bplane & 255) != 0) {
        set_error("b200_fq_gemm_tc_prepare_b: plane buffer too small or not 256-byte aligned (see b200_fq_gemm_tc_bplane_bytes)");
        return B200_ERR_WORKSPACE;
    }
    cudaStream_t st = (cudaStream_t)stream;
    int* flag = reinterpret_cast<int*>(bplane);
    B200_CUDA_TRY(cudaMemsetAsync(flag, 0, 4, st));
    return convert_b(B, sb_k, sb_n, K, N, round_up(K, TC_BK), reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(bplane) + kTcFlagBytes),
                     flag, st);
}

static int make_fq(const b200_act_quant* aq, FqParams* out, const char* who)
{
    *out = FqParams{nullptr, nullptr, 0.0f, 0};
    if (!aq) return B200_OK;
    if (!aq->maxval_dev || !aq->bias_dev || aq->n_bits < 2 || aq->sign_bits < 0 || aq->sign_bits > 1) {
        set_error(std::string(who) + ": bad b200_act_quant descriptor");
        return B200_ERR_ARGUMENT;
    }
    out->maxval = aq->maxval_dev;
    out->bias = aq->bias_dev;
    out->Mf = fminf(fmaxf(rintf(aq->mantissa_bits), 1.0f), (float)(aq->n_bits - aq->sign_bits));
    out->sign = aq->sign_bits;
    return B200_OK;
}

static int fq_gemm_tc_impl(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, const void* bplane, float* C,
                           int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bias_dev, void* workspace,
                           size_t workspace_bytes, void* stream, const b200_act_quant* aq_in = nullptr,
                           const b200_act_quant* aq_out = nullptr)
{
    NvtxRange nvtx_range("b200_fq_gemm_tc");
    if (int rc = check_device()) return rc;
    if (Mrows < 0 || K < 0 || N < 0) { set_error("b200_fq_gemm_tc: bad argument"); return B200_ERR_ARGUMENT; }
    if (Mrows == 0 || N == 0) return B200_OK;
    if ((K > 0 && (!A || !B)) || !C) { set_error("b200_fq_gemm_tc: NULL operand"); return B200_ERR_ARGUMENT; }
    cudaStream_t st = (cudaStream_t)stream;
    ASrc asrc = make_gemm_asrc(A, lda);
    if (int rc = apply_act_quant(&asrc, aq_in)) return rc;     // the fp32 fallback quantizes on load as well
    FqParams iq, oq;
    if (int rc = make_fq(aq_in, &iq, "b200_fq_gemm_tc_ex(input quantizer)")) return rc;
    if (int rc = make_fq(aq_out, &oq, "b200_fq_gemm_tc_ex(result quantizer)")) return rc;
    const BSrc bsrc{B, sb_k, sb_n};
    const PlainOp pop{bias_dev, oq, 0.0f, 0.0f, false};
    if (K == 0) {
        pair_gemm_kernel<PlainOp><<<pair_gemm_grid(Mrows, N), 256, 0, st>>>(asrc, bsrc, C, ldc, Mrows, K, N, pop, nullptr, 0);
        B200_LAUNCH_CHECK("pair_gemm_kernel<PlainOp>");
        set_path("fp32");
        return B200_OK;
    }
    if (workspace_bytes < b200_fq_gemm_tc_workspace_bytes(Mrows, K, N) || !workspace) {
        set_error("b200_fq_gemm_tc: workspace too small (see b200_fq_gemm_tc_workspace_bytes)");
        return B200_ERR_WORKSPACE;
    }
    EncodeTiledFn enc = nullptr;
    if (int rc = get_encode_fn(&enc)) return rc;
    const int64_t Kp = round_up(K, TC_BK);
    int* flag = reinterpret_cast<int*>(workspace);
    __nv_bfloat16* Ah = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(workspace) + kTcFlagBytes);
    const __nv_bfloat16* Bh = Ah + Mrows * Kp;
    if (bplane) {   // prepared right operand: its plane, and its "inexact" flag as the start value of this call's flag
        Bh = reinterpret_cast<const __nv_bfloat16*>(reinterpret_cast<const char*>(bplane) + kTcFlagBytes);
        B200_CUDA_TRY(cudaMemcpyAsync(flag, bplane, 4, cudaMemcpyDeviceToDevice, st));
    } else {
        B200_CUDA_TRY(cudaMemsetAsync(flag, 0, 4, st));
    }
    {
        const int64_t n = Mrows * (Kp / 8);
        to_bf16_kernel<true><<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(A, lda, 1, Ah, Mrows, K, Kp, flag, iq);
        B200_LAUNCH_CHECK("to_bf16_kernel<A>");
    }
    if (!bplane)
        if (int rc = convert_b(B, sb_k, sb_n, K, N, Kp, Ah + Mrows * Kp, flag, st)) return rc;

    const int nkb = (int)(Kp / TC_BK);
    static PerDeviceOnce once;
    int sms = 0;
    if (int rc = once.get(&sms, []() -> int {
            B200_CUDA_TRY(cudaFuncSetAttribute(fq_gemm_tc_kernel<256, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem<256, 4>::TOTAL));
            B200_CUDA_TRY(cudaFuncSetAttribute(fq_gemm_tc_kernel<128, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem<128, 4>::TOTAL));
            B200_CUDA_TRY(cudaFuncSetAttribute(fq_gemm_tc_kernel<192, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem<192, 4>::TOTAL));
            B200_CUDA_TRY(cudaFuncSetAttribute(fq_gemm_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Tc2Smem::TOTAL));
            return B200_OK;
        }))
        return rc;
    CUtensorMap mapA, mapB, mapC;
    if (int rc = make_map(enc, &mapA, Ah, Mrows, Kp, TC_BM)) return rc;
    // epilogue through TMA stores when C allows a tensor map (16-byte aligned base and pitch, extents in int32);
    // B200_TC_TMA_STORE=0: direct stores from registers
    const char* se = getenv("B200_TC_TMA_STORE");
    int tma_store = ((ldc & 3) == 0 && ((uintptr_t)C & 15) == 0 && Mrows < ((int64_t)1 << 31) && N < ((int64_t)1 << 31) &&
                     (bias_dev == nullptr || ((uintptr_t)bias_dev & 15) == 0) && !(se && atoi(se) == 0)) ? 1 : 0;
    memset(&mapC, 0, sizeof(mapC));
    if (tma_store)
        if (int rc = make_map_c(enc, &mapC, C, ldc, Mrows, N)) return rc;
    // CTA pairs (cta_group::2, 256 x 256 tiles) where there is enough work for them; B200_TC_PAIR=0: single-CTA kernel
    const char* pe = getenv("B200_TC_PAIR");
    const bool pair_ok = N > 128 && Mrows > 128 && sms >= 2 && !(pe && atoi(pe) == 0) &&
                         ((pe && atoi(pe) == 2) || ceil_div(Mrows, 2 * TC_BM) * ceil_div(N, TC2_BN) >= sms / 2);   // a full wave of pairs; 2: force (tests)
    if (pair_ok) {
        if (int rc = make_map(enc, &mapB, Bh, N, Kp, TC2_HALF_N)) return rc;
        const int m_blocks = (int)ceil_div(Mrows, 2 * TC_BM), n_tiles = m_blocks * (int)ceil_div(N, TC2_BN);
        const int pairs = std::min(n_tiles, sms / 2);
        fq_gemm_tc2_kernel<<<(unsigned)(2 * pairs), TC_THREADS, Tc2Smem::TOTAL, st>>>(
            mapA, mapB, mapC, tma_store, C, ldc, Mrows, N, nkb, m_blocks, n_tiles, bias_dev, flag, oq);
    } else {
        // single-CTA tiles 128 x BN: the BN that needs the fewest (waves x tile width) -- 4096 x 768 has 96 tiles of
        // 128 x 256 on 148 SMs but 128 tiles of 128 x 192
        auto waves_cost = [&](int bn) { return ceil_div(ceil_div(Mrows, TC_BM) * ceil_div(N, bn), sms) * bn; };
        int bn = N > 128 ? 256 : 128;
        if (N > 128 && waves_cost(192) < waves_cost(bn)) bn = 192;
        if (N > 128 && waves_cost(128) < waves_cost(bn)) bn = 128;
        if (const char* be = getenv("B200_TC_BN")) { const int v = atoi(be); if (v == 128 || v == 192 || v == 256) bn = v; }   // experiments
        auto launch1 = [&](auto bn_c) -> int {
            constexpr int BN = decltype(bn_c)::value, ST = 4;
            if (int rc = make_map(enc, &mapB, Bh, N, Kp, BN)) return rc;
            const int m_blocks = (int)ceil_div(Mrows, TC_BM), n_tiles = m_blocks * (int)ceil_div(N, BN);
            fq_gemm_tc_kernel<BN, ST><<<(unsigned)std::min(n_tiles, sms), TC_THREADS, TcSmem<BN, ST>::TOTAL, st>>>(
                mapA, mapB, mapC, tma_store, C, ldc, Mrows, N, nkb, m_blocks, n_tiles, bias_dev, flag, oq);
            return B200_OK;
        };
        int rc = bn == 256 ? launch1(std::integral_constant<int, 256>{})
                           : (bn == 192 ? launch1(std::integral_constant<int, 192>{}) : launch1(std::integral_constant<int, 128>{}));
        if (rc) return rc;
    }
    B200_LAUNCH_CHECK("fq_gemm_tc_kernel");
    // runs only if the convert pass found an operand that is not exact in bf16
    pair_gemm_kernel<PlainOp><<<pair_gemm_grid(Mrows, N), 256, 0, st>>>(asrc, bsrc, C, ldc, Mrows, K, N, pop, flag, 1);
    B200_LAUNCH_CHECK("pair_gemm_kernel<PlainOp>(conditional)");
    set_path(pair_ok ? (bplane ? "tcgen05-pair-preB" : "tcgen05-pair") : (bplane ? "tcgen05-preB" : "tcgen05"));
    return B200_OK;
}

extern "C" int b200_fq_gemm_tc(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, float* C,
                               int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bias_dev, void* workspace,
                               size_t workspace_bytes, void* stream)
{
    return fq_gemm_tc_impl(A, lda, B, sb_k, sb_n, nullptr, C, ldc, Mrows, K, N, bias_dev, workspace, workspace_bytes, stream);
}

extern "C" int b200_fq_gemm_tc_pre(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, const void* bplane,
                                   float* C, int64_t ldc, int64_t Mrows, int64_t K, int64_t N, const float* bias_dev,
                                   void* workspace, size_t workspace_bytes, void* stream)
{
    if (!bplane || K < 1) { set_error("b200_fq_gemm_tc_pre: needs a prepared plane (b200_fq_gemm_tc_prepare_b) and K >= 1"); return B200_ERR_ARGUMENT; }
    return fq_gemm_tc_impl(A, lda, B, sb_k, sb_n, bplane, C, ldc, Mrows, K, N, bias_dev, workspace, workspace_bytes, stream);
}

// The most general form (opt-in fusions): optional prepared right operand (bplane, may be NULL), optional input
// fake-quantizer applied while A is converted (aq_in: A then holds the RAW activations) and optional result
// fake-quantizer applied in the epilogue after the bias (aq_out).  b200_act_quant as in b200_approx_gemm_fbias_fq.
extern "C" int b200_fq_gemm_tc_ex(const float* A, int64_t lda, const float* B, int64_t sb_k, int64_t sb_n, const void* bplane,
                                  const b200_act_quant* aq_in, const b200_act_quant* aq_out, float* C, int64_t ldc,
                                  int64_t Mrows, int64_t K, int64_t N, const float* bias_dev, void* workspace,
                                  size_t workspace_bytes, void* stream)
{
    if (bplane && K < 1) { set_error("b200_fq_gemm_tc_ex: a prepared plane needs K >= 1"); return B200_ERR_ARGUMENT; }
    return fq_gemm_tc_impl(A, lda, B, sb_k, sb_n, bplane, C, ldc, Mrows, K, N, bias_dev, workspace, workspace_bytes, stream,
                           aq_in, aq_out);
}
