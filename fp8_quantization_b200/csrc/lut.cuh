// lut.cuh -- geometry of the pair table used by the table-driven approx GEMM (approx_gemm.cu).
//
// Main mode (with_s2nn2s_opt && quant_btw_mult_accu), operands on the FP8 grid:
//   a = sa * (1 + mA/2^M) * 2^LA ,  b = sb * (1 + mB/2^M) * 2^LB        (LA, LB = floor(log2|.|))
// The reference's addend (approx_matmul_whole_v9.py:29-110) is then a function of
//   t = LA + LB + bR , mA , mB , sb   and is odd in sa:
//   addend = sa * u * F(t, mA, mB, sb),   u = 2^(1 - bR - M)   (result sub-normal spacing)
// F is tabulated as bf16 (values have <= M+1 <= 6 significant bits).  The table index is ADDITIVE in
// the two operands so that the inner loop is  IADD -> LDS.U16 -> shift -> FFMA:
//   index = mB + 2^M*sb + 2^(M+1) * ( (tA_idx + tB_idx) + NT*mA )
//   tA_idx = LA - (1 - bA - M)            in [0, RA],  RA = 2^E + M
//   tB_idx = max(LB + bR - bA + 3, -RA) + RA   in [0, NTB-1]
//   t      = tA_idx + tB_idx - RA - M - 2
// Rows with tA_idx + tB_idx <= RA (t <= -M-2) are all zero (product rounds to 0), which is what
// makes the per-operand clamp of tB legal.  A zero `a` points at a separate all-zero region.
#pragma once
#include <stdint.h>

namespace b200 {

struct LutGeom {
    int E, M;
    int RA;          // max tA_idx
    int NTB;         // number of tB_idx values
    int NT;          // rows: RA + NTB
    int row_entries; // 2^(M+1): (sb, mB)
    int a_stride;    // entries per mA value: NT * row_entries
    int zeroA_off;   // entry offset of the all-zero region for a == 0
    int total_entries;
};

__host__ __device__ inline LutGeom make_lut_geom(int E, int M)
{
    LutGeom g;
    g.E = E;
    g.M = M;
    g.RA = (1 << E) + M;
    const int cneg = (E <= 2) ? 2 : 4;   // supports bA + bB[n] - bR >= -cneg on the table path
    g.NTB = g.RA + (1 << E) + 5 + cneg;
    g.NT = g.RA + g.NTB;
    g.row_entries = 1 << (M + 1);
    g.a_stride = g.NT * g.row_entries;
    g.zeroA_off = g.a_stride * (1 << M);
    g.total_entries = g.zeroA_off + g.NTB * g.row_entries;
    return g;
}

// Packed operand words (byte offsets into the bf16 table; bit 31 of an A word = sign of a).
constexpr uint32_t kSignBit = 0x80000000u;

}  // namespace b200
