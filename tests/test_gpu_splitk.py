"""Split-K of the table GEMM (csrc/approx_gemm.cu): layers whose output grid cannot fill the 148 SMs have
their k range cut into slices that are summed afterwards.  That changes the fp32 summation order, so the
kernel carries an exactness certificate and a gated unsplit fallback; the contract tested here is that the
output is BIT-IDENTICAL to the unsplit kernel's whether the certificate holds or not, and equal to the
oracle (approx/approx_matmul_whole_v9.py:10-169 restated) on a case the oracle finishes quickly."""
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


def grid_values(E, M, bias, shape, seed, top_frac=1.0):
    """Random signed values on the (E, M, bias) grid; `top_frac` < 1 keeps the codes in the lower part."""
    rng = np.random.default_rng(seed)
    n = int(2 ** (E + M) * top_frac)
    c = rng.integers(0, max(n, 2), size=shape).astype(np.int32)
    v = orc.decode(c >> M, c & (2 ** M - 1), E, M, bias)
    return (v * rng.choice([-1.0, 1.0], size=shape)).astype(np.float32)


def flags(fq):
    ws = fq.ops._WORKSPACES[torch.cuda.current_device()]
    return ws[:32].view(torch.int32).cpu().tolist()


def run(fq, a, b, bA, bB, bR, p, splits):
    if splits is None:
        os.environ.pop("B200_LUT_SPLITK", None)
    else:
        os.environ["B200_LUT_SPLITK"] = str(splits)
    try:
        out = fq.ops.approx_matmul(a, b, bA, bB, bR, p)
        return out, fq.ops.last_path(), flags(fq)
    finally:
        os.environ.pop("B200_LUT_SPLITK", None)


MAIN = dict(expo_width=4, mant_width=3, withComp=True, dnsmp_factor=3, with_approx=True, with_s2nn2s_opt=True,
            quant_btw_mult_accu=True)


@pytest.mark.parametrize("shape", [(392, 4608, 512), (200, 2304, 256), (49, 1000, 77)])
@pytest.mark.parametrize("top_frac,expect_exact", [(0.5, True), (1.0, False)])
def test_splitk_bit_identical_to_unsplit(fq, shape, top_frac, expect_exact):
    """Small operands: the certificate holds and the ordered slice sum is used.  Operands up to the format's
    maximum against a result grid that is far too fine: the certificate fails and the gated unsplit kernel
    recomputes the output.  Either way the bits equal the unsplit launch."""
    Mr, K, N = shape
    E, M = 4, 3
    bA, bB = 8, 8
    bR = 8 if expect_exact else 20      # a fine result grid makes the addends huge in units of u
    a = torch.from_numpy(grid_values(E, M, bA, (Mr, K), 1, top_frac)).cuda()
    b = torch.from_numpy(grid_values(E, M, bB, (K, N), 2, top_frac)).cuda()
    p = fq.ops.make_params(MAIN)
    ref, path1, _ = run(fq, a, b, bA, bB, bR, p, 1)
    assert path1 == "lut"
    for s in (2, 4, 7):
        got, path, fl = run(fq, a, b, bA, bB, bR, p, s)
        assert path == "lut-splitk", path
        assert fl[0] == 0
        assert (fl[6] == 0) == expect_exact, (fl, s)
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), (shape, s)


def test_splitk_default_choice_and_oracle(fq):
    """The cost model splits a ResNet-18 layer4-shaped GEMM at batch 8 (392 x 4608 x 512: 16 CTAs of 128 x 128
    on 148 SMs) without being told to; a sample of output columns is checked against the oracle."""
    Mr, K, N = 392, 4608, 512
    E, M, bA, bB, bR = 4, 3, 6, 10, 7
    a = grid_values(E, M, bA, (Mr, K), 3, 0.6)
    a[np.random.default_rng(4).random(a.shape) < 0.5] = 0.0           # post-ReLU activations
    b = grid_values(E, M, bB, (K, N), 5, 0.8)
    p = fq.ops.make_params(MAIN)
    got, path, fl = run(fq, torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), bA, bB, bR, p, None)
    assert path == "lut-splitk" and fl[6] == 0, (path, fl)
    T = orc.get_error_table_NN(E, M, True, 3)
    cols = [0, 1, 255, 511]
    rows = slice(0, 48)
    want = orc.approx_matmul(np.ascontiguousarray(a[rows]), np.ascontiguousarray(b[:, cols]), T, bA, bB, bR, E=E, M=M,
                             with_approx=True, with_s2nn2s_opt=True, quant_btw_mult_accu=True)
    g = got.cpu().numpy()[rows][:, cols]
    assert np.array_equal(g.view(np.uint32), np.asarray(want, np.float32).view(np.uint32))


def test_splitk_qamaa(fq):
    """quantize_after_mult_and_add on the table kernel: the result quantizer runs in the slice reduction."""
    Mr, K, N = 392, 2304, 256
    x = torch.from_numpy(grid_values(4, 3, 7, (Mr, K), 6, 0.7)).cuda()
    y = torch.from_numpy(grid_values(4, 3, 9, (K, N), 7, 0.7)).cuda()
    maxval = torch.tensor([float(x.abs().max() * y.abs().max() * 8)], device="cuda")
    kw = dict(x_bias=torch.tensor([7.0], device="cuda"), y_bias=torch.full((N,), 9.0, device="cuda"))
    outs = {}
    for s in (1, 3):
        os.environ["B200_LUT_SPLITK"] = str(s)
        try:
            outs[s] = fq.ops.qamaa_matmul(x, y, 8, maxval, 3.0, 1, **kw)
            path = fq.ops.last_path()
        finally:
            os.environ.pop("B200_LUT_SPLITK", None)
        assert path == ("qamaa-lut" if s == 1 else "qamaa-lut-splitk"), path
    assert torch.equal(outs[1].view(torch.int32), outs[3].view(torch.int32))
