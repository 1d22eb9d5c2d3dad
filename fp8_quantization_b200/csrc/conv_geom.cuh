// conv_geom.cuh -- conv geometry helpers shared by conv.cu and qamaa_gemm.cu.
// Reference: approx/approx_calculation.py:724-747 (im2col), :822-917 (run_forward).
#pragma once
#include "common.cuh"
#include "pair_gemm.cuh"

namespace b200 {

struct ConvDims {
    int64_t rows;   // B * Ho * Wo
    int Ho, Wo, Cg, Kg, Ng;
};

// Validates the geometry and derives output sizes; returns B200_OK or B200_ERR_ARGUMENT.
inline int conv_dims(const b200_conv_geom* g, ConvDims* d)
{
    if (!g || g->batch < 0 || g->C_in < 1 || g->C_out < 1 || g->kh < 1 || g->kw < 1 || g->stride_h < 1 ||
        g->stride_w < 1 || g->dil_h < 1 || g->dil_w < 1 || g->pad_h < 0 || g->pad_w < 0 || g->groups < 1 ||
        g->C_in % g->groups != 0 || g->C_out % g->groups != 0 || g->H < 1 || g->W < 1) {
        set_error("conv geometry: bad argument");
        return B200_ERR_ARGUMENT;
    }
    const int Ho = (g->H + 2 * g->pad_h - g->dil_h * (g->kh - 1) - 1) / g->stride_h + 1;
    const int Wo = (g->W + 2 * g->pad_w - g->dil_w * (g->kw - 1) - 1) / g->stride_w + 1;
    if (Ho < 1 || Wo < 1) { set_error("conv geometry: empty output"); return B200_ERR_ARGUMENT; }
    d->Ho = Ho;
    d->Wo = Wo;
    d->rows = g->batch * Ho * Wo;
    d->Cg = g->C_in / g->groups;
    d->Kg = d->Cg * g->kh * g->kw;
    d->Ng = g->C_out / g->groups;
    return B200_OK;
}

// Implicit-im2col view of group `grp` of the NCHW input.
inline ASrc make_conv_asrc(const float* x, const b200_conv_geom* g, const ConvDims& d, int grp)
{
    ASrc a{};
    a.p = x;
    a.lda = 0;
    a.conv = 1;
    a.C = g->C_in; a.H = g->H; a.W = g->W; a.kh = g->kh; a.kw = g->kw;
    a.sh = g->stride_h; a.sw = g->stride_w; a.ph = g->pad_h; a.pw = g->pad_w; a.dh = g->dil_h; a.dw = g->dil_w;
    a.Ho = d.Ho; a.Wo = d.Wo;
    a.c0 = grp * d.Cg;
    return a;
}

}  // namespace b200
