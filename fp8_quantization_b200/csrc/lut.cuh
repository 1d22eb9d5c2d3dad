// lut.cuh -- geometry of the pair table used by the table-driven approx GEMM (approx_gemm.cu).
//
// Main mode (with_s2nn2s_opt && quant_btw_mult_accu), operands on the FP8 grid:
//   a = sa * (1 + mA/2^M) * 2^LA ,  b = sb * (1 + mB/2^M) * 2^LB        (LA, LB = floor(log2|.|))
// The reference's addend (approx_matmul_whole_v9.py:29-110) is then a function of
//   t = LA + LB + bR , mA , mB , sa   and is odd in sb:
//   addend = sb * u * F(sa, t, mA, mB),   u = 2^(1 - bR - M)   (result sub-normal spacing)
// F is tabulated as bf16 (values have <= M+1 <= 6 significant bits).  The table index is ADDITIVE in
// the two operands so that the inner loop is  IADD -> LDS.U16 -> FHFMA.BF16 (fp32 += bf16 * bf16):
//   index = mB + 2^M * ( (tA_idx + tB_idx) + NT * (mA + 2^M * [a < 0]) )
//   tA_idx = LA - (1 - bA - M)            in [0, RA],  RA = 2^E + M
//   tB_idx = max(LB + bR - bA + 3, -RA) + RA   in [0, NTB-1]
//   t      = tA_idx + tB_idx - RA - M - 2
// Rows with tA_idx + tB_idx <= RA (t <= -M-2) are all zero (product rounds to 0), which is what
// makes the per-operand clamp of tB legal.  A zero `a` points at a separate all-zero region.
//
// Which operand carries what matters for shared-memory bank conflicts: in the GEMM all lanes of a
// warp share the A element and differ in the B element, so the B-dependent part of the index must
// span as few 4-byte words as possible.  It is (tB_idx, mB) only: 2^M bf16 per binade (16 B for
// M = 3).  The sign of `a` selects a table half (warp-uniform, costs nothing); the sign of `b`
// travels in the B word as a bf16 +-1.0 and is applied by the FHFMA multiplier.  (Round 1a had the
// sign of b in the index: twice the footprint, 1.94 wavefronts per lookup instead of 1.18 --
// profiles/r1a_approx_gemm_lut_full.txt.)
#pragma once
#include <stdint.h>

namespace b200 {

struct LutGeom {
    int E, M;
    int RA;          // max tA_idx
    int NTB;         // number of tB_idx values
    int NT;          // rows: RA + NTB
    int row_entries; // 2^M: mB
    int a_stride;    // entries per (sa, mA) value: NT * row_entries
    int zeroA_off;   // entry offset of the all-zero region for a == 0
    int total_entries;
    int sidx;        // 1: "signed index" layout for tables that are not odd in the operand signs
    int s_stride;    // entries per sign region: a_stride * 2^M
    int v5;          // 1: table of the v5 integer-add GEMM (ofuf_mode 1), indexed by code exponents
    int v5_lo;       // v5: most negative supported eB - (bA + bB - bR)
};

// Signed-index layout (sidx): the addend depends on the signs only through s = [a<0] xor [b<0], but for
// a few cells at the round-to-zero edge of one table (E2M5, withComp, dnsmp 3) it is NOT odd in s: the
// reference takes the sign from a golden product that already rounded to zero
// (approx_matmul_whole_v9.py:68 after :36).  There the sign of b cannot ride on the FHFMA multiplier.
// Both signs go into the index instead, additively: [a<0]*S + [b<0]*S selects one of THREE regions
// (s = 0, s = 1, and a second copy of s = 0 for 1+1), each [mA][t][mB]; the multiplier is +1.0.
__host__ __device__ inline LutGeom make_lut_geom(int E, int M, bool sidx = false)
{
    LutGeom g;
    g.E = E;
    g.M = M;
    g.RA = (1 << E) + M;
    const int cneg = (E <= 2) ? 2 : 4;   // supports bA + bB[n] - bR >= -cneg on the table path
    g.NTB = g.RA + (1 << E) + 5 + cneg;
    g.NT = g.RA + g.NTB;
    g.row_entries = 1 << M;
    g.a_stride = g.NT * g.row_entries;
    g.sidx = sidx ? 1 : 0;
    g.s_stride = g.a_stride << M;
    g.zeroA_off = g.s_stride * (sidx ? 3 : 2);
    g.total_entries = g.zeroA_off + g.NTB * g.row_entries;
    g.v5 = 0;
    g.v5_lo = 0;
    return g;
}

// v5 flavour: same entry layout [sa][mA][row][mB], row = tA_idx + tB_idx with
//   tA_idx = eA                               in [0, RA],  RA = 2^E - 1   (code exponent of a)
//   tB_idx = eB - (bA + bB - bR) + lo + RA + 1   in [RA + 1, NTB - 1]      (code exponent of b, re-biased per column)
//   row - (lo + RA + 1) = eA + eB - (bA + bB - bR), the exponent part of the product code t (v5 :159)
// tB_idx = 0 is reserved for padding: rows <= RA are all zero.  Exponent sums outside [-lo, 2^E + 3] + [0, RA]
// (never met with calibrated biases) raise the off-grid flag instead.
__host__ __device__ inline LutGeom make_lut_geom_v5(int E, int M)
{
    LutGeom g;
    g.E = E;
    g.M = M;
    g.RA = (1 << E) - 1;
    g.v5_lo = (1 << E) + 8;
    g.NTB = g.RA + 1 + g.v5_lo + (1 << E) + 4;
    g.NT = g.RA + g.NTB;
    g.row_entries = 1 << M;
    g.a_stride = g.NT * g.row_entries;
    g.sidx = 0;
    g.s_stride = g.a_stride << M;
    g.zeroA_off = g.s_stride * 2;
    g.total_entries = g.zeroA_off + g.NTB * g.row_entries;
    g.v5 = 1;
    return g;
}

// Packed operand words.  A word: byte offset of the (sa, tA_idx, mA) row start in the bf16 table.
// B word: low 16 bits = byte offset 2 * (mB + 2^M * tB_idx)  (<= 2 * 2^M * NTB < 2^16),
//         high 16 bits = bf16 +1.0 / -1.0 (sign of b), the FHFMA multiplier.
constexpr uint32_t kBf16One = 0x3F80u, kBf16MinusOne = 0xBF80u;

}  // namespace b200
