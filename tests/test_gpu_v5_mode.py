"""SURVEY.md 8(a) row A7 inside the GEMM and the conv: ofuf_mode = 1, the v5 integer-add GEMM in which
sim_hw_add_OFUF / with_OF_opt / with_UF_opt are LIVE (approx/approx_matmul_whole_v5.py:10-182; approx_v9 ignores
them, and so does the default ofuf_mode = 0).

  * tests/golden/v5_gemm.npz: outputs of the UNMODIFIED reference's v5 custom_matmul_vectorize on K > 1 GEMMs, 3
    formats x 2 biases x {comp table, zeros} x all 8 flag combinations, operands on and off the grid -- table kernel
    and arithmetic kernel.  Tolerance: bit exact whenever the fp32 sum is order independent (always with
    sim_hw_add_OFUF: every addend is then a multiple of u below 2^(2^E+M)); 1e-6 norm-wise otherwise.
  * per-column weight biases, ragged shapes, K tails: against the oracle (bit exact: same k-ascending fp32 order).
  * conv form (groups 1, grouped, depthwise) against the oracle's im2col + v5 GEMM: padding taps are real zeros of
    the reference's im2col matrix and take part in v5's integer add.
"""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def fq():
    import fp8_quantization_b200 as pkg
    pkg._lib.load()
    assert torch.cuda.is_available()
    return pkg


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0))))


def params(fq, E, M, wc, sim, of, uf, **kw):
    return fq.ops.make_params(dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=bool(wc), sim_hw_add_OFUF=bool(sim),
                                   with_OF_opt=bool(of), with_UF_opt=bool(uf)), ofuf_mode="v5", **kw)


@pytest.mark.parametrize("force_generic", [False, True])
def test_v5_gemm_fixture(fq, force_generic):
    z = np.load(os.path.join(GOLD, "v5_gemm.npz"))
    bad, nexact, paths = [], 0, set()
    for i, (E, M, bias, ongrid, wc, sim, of, uf, g) in enumerate(z["cfgs"].tolist()):
        A, B, want = z[f"A{g}"], z[f"B{g}"], z[f"r{i}"]
        got = fq.ops.approx_matmul(dev(A), dev(B), bias, bias, bias,
                                   params(fq, E, M, wc, sim, of, uf, force_generic=force_generic)).cpu().numpy()
        paths.add(fq.ops.last_path())
        if same_bits(got, want):
            nexact += 1
            continue
        if sim:
            bad.append((i, "not bit exact with sim_hw_add_OFUF", E, M, bias, ongrid, wc, of, uf))
            continue
        absA, absB = np.abs(A).astype(np.float64), np.abs(B).astype(np.float64)
        T = orc.v5_comp_table(E, M, bool(wc), 3)
        mag = np.abs(orc.v5_matmul(absA, absB, T, E, M, bias, bias, bias, sim, of, uf, sum_mode=1)).astype(np.float64)
        if np.any(np.abs(got.astype(np.float64) - want) > 1e-6 * mag + 1e-30):
            bad.append((i, "beyond 1e-6 norm-wise", E, M, bias, ongrid, wc, sim, of, uf))
    assert not bad, bad[:6]
    assert paths == ({"generic-v5"} if force_generic else {"lut-v5"}), paths
    assert nexact >= len(z["cfgs"]) * 3 // 4


def grid_values(rng, E, M, bias, shape, zero_frac=0.2):
    expo = rng.integers(0, 2 ** E, size=shape)
    mant = rng.integers(0, 2 ** M, size=shape)
    v = np.where(expo == 0, 2.0 ** (1 - bias) * mant / 2 ** M, 2.0 ** (expo - bias) * (1 + mant / 2 ** M))
    v = v * rng.choice([-1.0, 1.0], size=shape)
    v[rng.random(shape) < zero_frac] = 0.0
    return v.astype(np.float32)


@pytest.mark.parametrize("fmt", [(4, 3), (3, 4), (2, 5)])
@pytest.mark.parametrize("flags", [(0, 0, 0), (1, 0, 0), (1, 1, 1), (1, 0, 1)])
def test_v5_gemm_per_column_bias_vs_oracle(fq, fmt, flags):
    E, M = fmt
    sim, of, uf = flags
    rng = np.random.default_rng(1000 * E + 10 * sim + of + uf)
    Mr, K, N = 300, 203, 130
    bA, bR = 2 ** (E - 1) + 1, 2 ** (E - 1) + 2
    bB = rng.integers(2 ** (E - 1), 2 ** (E - 1) + 5, size=N).astype(np.int32)
    A = grid_values(rng, E, M, bA, (Mr, K))
    B = np.stack([grid_values(rng, E, M, int(b), (K,)) for b in bB], axis=1)
    T = orc.v5_comp_table(E, M, True, 3)
    want = orc.v5_matmul(A, B, T, E, M, bA, bB, bR, sim, of, uf)
    for force_generic in (False, True):
        p = params(fq, E, M, 1, sim, of, uf, force_generic=force_generic)
        got = fq.ops.approx_matmul(dev(A), dev(B), bA, dev(bB), bR, p).cpu().numpy()
        assert same_bits(got, want), (fmt, flags, force_generic, fq.ops.last_path())
    # the default mode ignores the three flags exactly like approx_v9
    base = dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True, with_s2nn2s_opt=True,
                quant_btw_mult_accu=True)
    y0 = fq.ops.approx_matmul(dev(A), dev(B), bA, dev(bB), bR, fq.ops.make_params(base))
    y1 = fq.ops.approx_matmul(dev(A), dev(B), bA, dev(bB), bR, fq.ops.make_params(
        dict(base, sim_hw_add_OFUF=True, with_OF_opt=True, with_UF_opt=True)))
    assert torch.equal(y0, y1)


@pytest.mark.parametrize("geom", [(2, 8, 9, 9, 12, 3, 1, 1, 1), (2, 8, 10, 7, 8, 3, 2, 1, 2), (2, 6, 9, 9, 6, 3, 1, 1, 6)])
def test_v5_conv_vs_oracle(fq, geom):
    """(B, C, H, W, C_out, k, stride, pad, groups); depthwise groups take the arithmetic all-groups kernel."""
    Bn, C, H, W, Co, k, st, pad, groups = geom
    E, M = 4, 3
    rng = np.random.default_rng(sum(geom))
    bA, bR = 9, 10
    bW = rng.integers(8, 12, size=Co).astype(np.int32)
    x = grid_values(rng, E, M, bA, (Bn, C, H, W), zero_frac=0.4)
    w = np.stack([grid_values(rng, E, M, int(b), (C // groups, k, k), zero_frac=0.05) for b in bW])
    col, Ho, Wo = orc.im2col(x, k, k, (st, st), (pad, pad), (1, 1))
    T = orc.v5_comp_table(E, M, True, 3)
    Kg, Ng = (C // groups) * k * k, Co // groups
    want = np.zeros((col.shape[0], Co), np.float32)
    for g in range(groups):
        want[:, g * Ng:(g + 1) * Ng] = orc.v5_matmul(col[:, g * Kg:(g + 1) * Kg], w[g * Ng:(g + 1) * Ng].reshape(Ng, Kg).T.copy(),
                                                      T, E, M, bA, bW[g * Ng:(g + 1) * Ng], bR, 1, 1, 1)
    p = params(fq, E, M, 1, 1, 1, 1, single_column_quirk=(Ng == 1))
    got = fq.ops.approx_conv2d(dev(x), dev(w), torch.tensor([float(bA)]), dev(bW.astype(np.float32)), torch.tensor([float(bR)]),
                               p, (st, st), (pad, pad), (1, 1), groups).cpu().numpy()
    assert same_bits(got, want), (geom, fq.ops.last_path())
    assert fq.ops.last_path() == ("single-column-v5" if Ng == 1 else "lut-v5")


def test_v5_mode_through_the_operator_class(fq):
    """Opt-in from the layer: an extra `ofuf_mode` key in custom_approx_params (not one of the reference's 13 keys)."""
    from fp8_quantization_b200.approx_calculation import QCustomLinearTorch
    outs = {}
    for mode in (None, "v5"):
        cap = dict(expo_width=4, mant_width=3, dnsmp_factor=3, withComp=True, with_approx=True, with_s2nn2s_opt=True,
                   sim_hw_add_OFUF=True, with_OF_opt=True, with_UF_opt=True, golden_clip_OF=False,
                   quant_btw_mult_accu=True, debug_mode=False, self_check_mode=False)
        if mode:
            cap["ofuf_mode"] = mode
        torch.manual_seed(5)
        layer = QCustomLinearTorch(96, 40, bias=True, method="fp_quantizer", act_method="fp_quantizer", n_bits=8,
                                   n_bits_act=8, per_channel_weights=True, weight_range_method="current_minmax",
                                   act_range_method="allminmax", quantize_input=True,
                                   fp8_kwargs=dict(mantissa_bits=3, set_maxval=True), custom_approx_params=cap,
                                   run_method=dict(approx_flag=True, quantize_after_mult_and_add=False,
                                                   res_quantizer_flag=True, original_quantize_res=False)).cuda().eval()
        x = torch.randn(24, 96, generator=torch.Generator().manual_seed(6)).cuda()
        with torch.inference_mode():
            layer.quantized()
            layer.estimate_ranges()
            layer(x)
            layer.fix_ranges()
            outs[mode] = layer(x)
            if mode == "v5":
                assert fq.ops.last_path() == "lut-v5"
                xq = layer.activation_quantizer(x)
                wq, bias = layer.get_params()
                want = orc.v5_matmul(xq.cpu().numpy(), wq.t().contiguous().cpu().numpy(), orc.v5_comp_table(4, 3, True, 3),
                                     4, 3, int(layer.get_acts_fp_bias().item()),
                                     layer.get_weights_fp_bias().reshape(-1).to(torch.int32).cpu().numpy(),
                                     int(layer.get_res_fp_bias().item()), 1, 1, 1)
                assert same_bits(outs[mode].cpu().numpy(), want + bias.cpu().numpy())
    assert not torch.equal(outs[None], outs["v5"])
