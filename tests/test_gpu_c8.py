"""Byte-cell variant of the table GEMM (csrc/lut_gemm.cuh, C8 instantiations in csrc/approx_gemm_c8.cu): for
mant_width >= 4 the pair table is also kept with int8 cells (half the shared-memory footprint a warp's lookups
spread over) and sums are accumulated as integers.  Cells beyond +-127 are zero in that table; the A elements whose
table rows could meet one ("escapes") are listed by the predecode pass and a correction pass adds the dropped 16-bit
cells.  Both variants are enqueued on the same operand planes and the choice is made on the device; the contract
tested here is that the output is BIT-IDENTICAL to the 16-bit-cell kernel's (params.force_table16) in every
situation -- escapes rare and common, gate closed, split-K, stream-K, sparse and dense activations, the conv forms --
and equal to the oracle (approx/approx_matmul_whole_v9.py:10-169 restated) on a column sample."""
import os

import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fq():
    import fp8_quantization_b200 as pkg
    pkg._lib.load()
    return pkg


def approx_dict(M, withComp=True):
    return dict(expo_width=7 - M, mant_width=M, dnsmp_factor=3, withComp=withComp, with_approx=True,
                with_s2nn2s_opt=True, quant_btw_mult_accu=True)


def operands(fq, M, Mr, K, N, seed, relu=False, res_scale=1.5, x_scale=0.5):
    """Fake-quantized Gaussian operands and the three exponent biases, as bench.py's sweep makes them."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(Mr, K, generator=g) * x_scale
    if relu:
        x = x.clamp_min(0.0)
    w = torch.randn(N, K, generator=g) * 0.02
    x, w = x.cuda(), w.cuda()
    mant = torch.tensor([float(M)])
    xq, bA = fq.ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
    wq, bB = fq.ops.fp8_fakequant(w, 8, w.abs().amax(dim=1), mant, 1)
    mv = ((xq[:256] @ wq.t()).abs().max() * res_scale).reshape(1)
    _, bR = fq.ops.fp8_fakequant(mv.clone(), 8, mv, mant, 1)
    return xq, wq, bA, bB, bR, mv


def both(fq, M, xq, wq, bA, bB, bR, withComp=True, env=None):
    """(byte-cell-eligible call, forced 16-bit call, variant record of the first)."""
    env = env or {}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        got = fq.ops.approx_matmul(xq, wq.t(), bA, bB, bR, fq.ops.make_params(approx_dict(M, withComp)))
        path = fq.ops.last_path()
        var = fq.ops.table_variant()
        want = fq.ops.approx_matmul(xq, wq.t(), bA, bB, bR, fq.ops.make_params(approx_dict(M, withComp), force_table16=True))
        var16 = fq.ops.table_variant()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    assert var16["variant"] == "16-bit"
    return got, want, var, path


def oracle_columns(M, xq, wq, bA, bB, bR, cols, withComp=True):
    T = orc.get_error_table_NN(7 - M, M, withComp, 3)
    B = np.ascontiguousarray(wq[cols].cpu().numpy().T)
    bBc = bB.reshape(-1)[cols].to(torch.int32).cpu().numpy()
    return orc.approx_matmul(xq.cpu().numpy(), B, T, int(bA.item()), bBc, int(bR.item()), E=7 - M, M=M, with_approx=1,
                             with_s2nn2s_opt=1, quant_btw_mult_accu=1)


@pytest.mark.parametrize("M", [4, 5])
@pytest.mark.parametrize("withComp", [True, False])
@pytest.mark.parametrize("shape", [(300, 520, 200), (1024, 1024, 384), (788, 768, 768)])
@pytest.mark.parametrize("relu", [False, True])
def test_byte_cells_equal_16bit_cells(fq, M, withComp, shape, relu):
    Mr, K, N = shape
    xq, wq, bA, bB, bR, _ = operands(fq, M, Mr, K, N, seed=M * 100 + Mr + (7 if relu else 0), relu=relu)
    got, want, var, path = both(fq, M, xq, wq, bA, bB, bR, withComp)
    assert path.startswith("lut"), path
    # short K: the result range sits close to the single products, whole binades of A are escapes and the gate may
    # close (the bits must not care); K = 1024 is comfortably inside
    if K >= 1024:
        assert var["variant"] == "byte-cells", var
    assert torch.equal(got.view(torch.int32), want.view(torch.int32)), var
    cols = [0, 1, N // 2, N - 1]
    ref = oracle_columns(M, xq, wq, bA, bB, bR, cols, withComp)
    assert np.array_equal(got[:, cols].cpu().numpy().view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("M", [4, 5])
@pytest.mark.parametrize("sparse", ["0", "1"])
def test_both_k_loops(fq, M, sparse):
    """B200_LUT_SPARSE forces the predicated (0) / the branchy (1) k loop whatever the zero statistics say."""
    xq, wq, bA, bB, bR, _ = operands(fq, M, 640, 1024, 256, seed=5, relu=True)
    got, want, var, _ = both(fq, M, xq, wq, bA, bB, bR, env={"B200_LUT_SPARSE": sparse})
    assert var["variant"] == "byte-cells", var
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))


@pytest.mark.parametrize("M", [4, 5])
def test_escapes_take_the_16bit_cell(fq, M):
    """Shrinking the result range makes the grid finer and the cells larger in units of u, which lowers the last
    all-int8 table row: first only the operands' top binade (thinly populated, plus 40 injected maxima) lies above
    it.  Those A elements are escapes: listed, and the cells the byte table dropped for them come from the full table."""
    hit = None
    for res_scale in (1.5, 0.75, 0.38, 0.19, 0.09, 0.045):
        xq, wq, bA, bB, bR, _ = operands(fq, M, 512, 768, 256, seed=11, res_scale=res_scale)
        top = xq.abs().max()
        rng = np.random.default_rng(3)
        ii, jj = rng.integers(0, 512, 40), rng.integers(0, 768, 40)
        xq[torch.from_numpy(ii).cuda(), torch.from_numpy(jj).cuda()] = top * torch.from_numpy(rng.choice([-1.0, 1.0], 40)).float().cuda()
        got, want, var, _ = both(fq, M, xq, wq, bA, bB, bR)
        assert var["offgrid"] == 0, var
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), (res_scale, var)
        if var["variant"] == "byte-cells" and var["escapes"] > 0:
            hit = (xq, wq, bA, bB, bR, got, var)
            break
    assert hit is not None, f"no configuration exercised the escape path: {var}"
    xq, wq, bA, bB, bR, got, var = hit
    assert var["escapes"] >= 40
    cols = [3, 100]
    ref = oracle_columns(M, xq, wq, bA, bB, bR, cols)
    assert np.array_equal(got[:, cols].cpu().numpy().view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("M", [4, 5])
def test_gate_closes_when_escapes_are_common_or_sums_inexact(fq, M):
    """A much finer result grid: more than 1/8 of the A elements lie above the last all-int8 row (or the integer sum is
    no longer provably the fp32 sum) -- the device picks the 16-bit kernel; the bits are the same either way."""
    xq, wq, bA, bB, bR, _ = operands(fq, M, 512, 768, 256, seed=12)
    got, want, var, _ = both(fq, M, xq, wq, bA, bB, bR + 4)
    assert var["offgrid"] == 0, var
    assert var["variant"] == "16-bit", var
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))


@pytest.mark.parametrize("M", [4, 5])
def test_rare_top_binade_of_b(fq, M):
    """The channel maxima of the weights sit alone in B's top binade.  Measured against them a whole populated binade
    of A would be escapes; the pick kernel declares them rare instead, the A pass compares against the next level, and
    the correction pass covers (rare B) x (non-escape A).  Result grids one and two binades finer than calibrated make
    those pairs really exceed int8."""
    seen = []
    hit = False
    for step in (0, 1, 2):
        xq, wq, bA, bB, bR, _ = operands(fq, M, 400, 768, 300, seed=21)
        got, want, var, _ = both(fq, M, xq, wq, bA, bB, bR + step)
        seen.append(var)
        assert var["offgrid"] == 0, var
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), var
        if var["variant"] == "byte-cells" and var["rare_b"] > 0:
            assert var["tb_eff"] < var["tmax"][1], var
            cols = [0, 150, 299]
            ref = oracle_columns(M, xq, wq, bA, bB, bR + step, cols)
            assert np.array_equal(got[:, cols].cpu().numpy().view(np.uint32), ref.view(np.uint32))
            hit = True
    assert hit or M == 5, f"the rare-B class never engaged: {seen}"   # (E2M5 withComp: signed-index table, no rare class)


@pytest.mark.parametrize("M", [4, 5])
def test_split_k_and_stream_k(fq, M):
    xq, wq, bA, bB, bR, _ = operands(fq, M, 392, 2304, 256, seed=13, relu=True)
    ref, _, _, _ = both(fq, M, xq, wq, bA, bB, bR, env={"B200_LUT_SPLITK": "1"})
    for env, name in (({"B200_LUT_SPLITK": "2"}, "lut-splitk"), ({"B200_LUT_SPLITK": "4"}, "lut-splitk"),
                      ({"B200_LUT_STREAMK": "1"}, "lut-streamk")):
        got, want, var, path = both(fq, M, xq, wq, bA, bB, bR, env=env)
        if M == 4:
            assert path == name, (path, env)
        assert var["variant"] == "byte-cells", var
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), env
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), env


@pytest.mark.parametrize("M", [4, 5])
def test_quantize_after_mult_and_add(fq, M):
    """qamaa rebuilds both tables per call from the result quantizer's maxval; the arithmetic kernel is the check."""
    xq, wq, bA, bB, _, mv = operands(fq, M, 520, 640, 200, seed=14)
    mant = torch.tensor([float(M)])
    got = fq.ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1, x_bias=bA, y_bias=bB)
    path, var = fq.ops.last_path(), fq.ops.table_variant()
    os.environ["B200_QAMAA_GENERIC"] = "1"
    try:
        want = fq.ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1, x_bias=bA, y_bias=bB)
        assert fq.ops.last_path() == "qamaa-generic"
    finally:
        os.environ.pop("B200_QAMAA_GENERIC", None)
    assert path.startswith("qamaa-lut"), path
    assert var["variant"] == "byte-cells", var
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))


@pytest.mark.parametrize("M", [4, 5])
@pytest.mark.parametrize("conv", [dict(cin=64, cout=96, k=1, pad=0, hw=14), dict(cin=24, cout=320, k=3, pad=1, hw=12)])
def test_conv_forms(fq, M, conv):
    """Pointwise convs (elementwise predecode) and the im2col-plane predecode flag escapes too; the conv gather
    (N <= 256, 3x3) keeps the 16-bit kernel."""
    g = torch.Generator().manual_seed(15)
    x = (torch.randn(4, conv["cin"], conv["hw"], conv["hw"], generator=g) * 0.5).cuda()
    w = (torch.randn(conv["cout"], conv["cin"], conv["k"], conv["k"], generator=g) * 0.05).cuda()
    mant = torch.tensor([float(M)])
    xq, bA = fq.ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
    wq, bB = fq.ops.fp8_fakequant(w.reshape(conv["cout"], -1), 8, w.abs().amax(dim=(1, 2, 3)), mant, 1)
    wq = wq.reshape(w.shape)
    y = torch.nn.functional.conv2d(xq, wq, padding=conv["pad"])
    mv = (y.abs().max() * 8.0).reshape(1)      # short K: a roomy result range keeps the products' cells within int8
    _, bR = fq.ops.fp8_fakequant(mv.clone(), 8, mv, mant, 1)
    args = ((1, 1), (conv["pad"], conv["pad"]), (1, 1), 1)
    got = fq.ops.approx_conv2d(xq, wq, bA, bB.reshape(-1), bR, fq.ops.make_params(approx_dict(M)), *args)
    var = fq.ops.table_variant()
    want = fq.ops.approx_conv2d(xq, wq, bA, bB.reshape(-1), bR, fq.ops.make_params(approx_dict(M), force_table16=True), *args)
    assert var["variant"] == "byte-cells", var
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))
