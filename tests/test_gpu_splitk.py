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
    dev = torch.cuda.current_device()
    ws = fq.ops._WORKSPACES[(dev, torch.cuda.current_stream(dev).cuda_stream)]
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
    """A ResNet-18 layer4-shaped GEMM at batch 8 (392 x 4608 x 512: 16 CTAs of 128 x 128 on 148 SMs) is split
    along K (split-K or stream-K) without being told to; a sample of output columns is checked against the oracle."""
    Mr, K, N = 392, 4608, 512
    E, M, bA, bB, bR = 4, 3, 6, 10, 7
    a = grid_values(E, M, bA, (Mr, K), 3, 0.6)
    a[np.random.default_rng(4).random(a.shape) < 0.5] = 0.0           # post-ReLU activations
    b = grid_values(E, M, bB, (K, N), 5, 0.8)
    p = fq.ops.make_params(MAIN)
    got, path, fl = run(fq, torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda(), bA, bB, bR, p, None)
    assert path in ("lut-splitk", "lut-streamk") and fl[6] == 0, (path, fl)     # split or stream-K, whichever measured faster
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


CONV_CASES = [
    # B, C, H, W, C_out, k, stride, pad, dil, groups
    (2, 16, 14, 14, 32, 3, 1, 1, 1, 1),
    (2, 16, 14, 14, 32, 3, 2, 1, 1, 1),
    (3, 8, 12, 10, 24, 3, 1, 2, 2, 1),       # dilation, non-square
    (2, 3, 32, 32, 16, 7, 2, 3, 1, 1),       # the ResNet stem shape
    (2, 32, 8, 8, 64, 1, 2, 0, 1, 1),        # 1x1 stride-2 downsample
    (2, 32, 8, 8, 64, 1, 1, 0, 1, 1),        # pointwise
    (2, 16, 10, 10, 32, 3, 1, 1, 1, 4),      # grouped, 8 columns per group
    (4, 64, 7, 7, 128, 3, 1, 1, 1, 1),       # 49-pixel images: image size not a multiple of 4 words per channel
    (1, 5, 7, 9, 6, 3, 1, 1, 1, 1),          # image of 315 words: the gather path must decline (alignment)
    (3, 24, 6, 6, 40, 1, 1, 0, 1, 1),        # pointwise, 108 rows / K = 24: row and k padding of the 1x1 decode pass
    (2, 20, 10, 10, 24, 1, 1, 0, 1, 2),      # pointwise, grouped
    (2, 12, 5, 5, 16, 1, 1, 0, 1, 1),        # pointwise on 25-pixel images: the 1x1 pass must decline (H*W % 4)
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("mode", ["approx", "qamaa"])
def test_conv_gather_equals_im2col_plane(fq, case, mode):
    """Conv form: decoding the NCHW input once and gathering in the GEMM's tile loader must give the
    bits of the older path that writes the im2col plane (B200_CONV_GATHER=0), which the layer fixtures from the
    reference's QCustomBNConv2dTorch pinned (approx/approx_calculation.py:724-747, :846-906)."""
    B, C, H, W, Co, k, st, pad, dil, groups = case
    E, M, bA, bW = 4, 3, 7, 9
    x = grid_values(E, M, bA, (B, C, H, W), 11, 0.8)
    x[np.random.default_rng(12).random(x.shape) < 0.4] = 0.0
    w = grid_values(E, M, bW, (Co, C // groups, k, k), 13, 0.8)
    xd, wd = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda()
    xb = torch.tensor([float(bA)], device="cuda")
    wb = torch.full((Co,), float(bW), device="cuda")
    outs = []
    for env in ("0", "2"):      # never / whenever the geometry allows it
        os.environ["B200_CONV_GATHER"] = env
        try:
            if mode == "approx":
                rb = torch.tensor([8.0], device="cuda")
                o = fq.ops.approx_conv2d(xd, wd, xb, wb, rb, fq.ops.make_params(MAIN), (st, st), (pad, pad), (dil, dil), groups)
            else:
                mv = torch.tensor([float(np.abs(x).max() * np.abs(w).max() * 16)], device="cuda")
                o = fq.ops.qamaa_conv2d(xd, wd, 8, mv, 3.0, 1, (st, st), (pad, pad), (dil, dil), groups, x_bias=xb, w_bias=wb)
            assert fq.ops.last_path().startswith("lut" if mode == "approx" else "qamaa-lut"), fq.ops.last_path()
            outs.append(o)
        finally:
            os.environ.pop("B200_CONV_GATHER", None)
    assert float(outs[0].abs().max()) > 0
    assert torch.equal(outs[0].view(torch.int32), outs[1].view(torch.int32)), case


def _streamk(fn):
    os.environ["B200_LUT_STREAMK"] = "1"
    try:
        return fn(), __import__("fp8_quantization_b200").ops.last_path()
    finally:
        os.environ.pop("B200_LUT_STREAMK", None)


@pytest.mark.parametrize("shape", [(3152, 768, 768), (392, 4608, 512), (1000, 200, 130), (300, 1152, 128), (129, 96, 257)])
@pytest.mark.parametrize("top_frac,expect_exact", [(0.5, True), (1.0, False)])
def test_streamk_bit_identical_to_unsplit(fq, shape, top_frac, expect_exact):
    """Stream-K (persistent CTAs over equal ranges of (tile, k-tile) units, partial tiles added in k order by the
    fix-up kernel): bit-identical to the unsplit launch whether the exactness certificate holds or fails."""
    Mr, K, N = shape
    E, M = 4, 3
    bA, bB = 8, 8
    bR = 8 if expect_exact else 20
    a = torch.from_numpy(grid_values(E, M, bA, (Mr, K), 21, top_frac)).cuda()
    a[torch.rand(a.shape, device="cuda") < 0.3] = 0.0
    b = torch.from_numpy(grid_values(E, M, bB, (K, N), 22, top_frac)).cuda()
    p = fq.ops.make_params(MAIN)
    ref, path1, _ = run(fq, a, b, bA, bB, bR, p, 1)
    assert path1 == "lut"
    got, path = _streamk(lambda: fq.ops.approx_matmul(a, b, bA, bB, bR, p))
    fl = flags(fq)
    assert path == "lut-streamk", path
    assert fl[0] == 0 and (fl[6] == 0) == expect_exact, fl
    assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), shape


def test_streamk_qamaa_and_conv(fq):
    """Stream-K under the quantizing epilogue (the result quantizer runs in the fix-up) and with the conv gather
    loader (the row setup is redone per tile)."""
    Mr, K, N = 392, 2304, 256
    x = torch.from_numpy(grid_values(4, 3, 7, (Mr, K), 6, 0.7)).cuda()
    y = torch.from_numpy(grid_values(4, 3, 9, (K, N), 7, 0.7)).cuda()
    maxval = torch.tensor([float(x.abs().max() * y.abs().max() * 8)], device="cuda")
    kw = dict(x_bias=torch.tensor([7.0], device="cuda"), y_bias=torch.full((N,), 9.0, device="cuda"))
    os.environ["B200_LUT_SPLITK"] = "1"
    try:
        ref = fq.ops.qamaa_matmul(x, y, 8, maxval, 3.0, 1, **kw)
    finally:
        os.environ.pop("B200_LUT_SPLITK", None)
    got, path = _streamk(lambda: fq.ops.qamaa_matmul(x, y, 8, maxval, 3.0, 1, **kw))
    assert path == "qamaa-lut-streamk", path
    assert torch.equal(got.view(torch.int32), ref.view(torch.int32))
    for case in CONV_CASES[:5] + CONV_CASES[6:8]:
        B, C, H, W, Co, k, st, pad, dil, groups = case
        E, M, bA, bW = 4, 3, 7, 9
        xn = grid_values(E, M, bA, (B, C, H, W), 11, 0.8)
        xn[np.random.default_rng(12).random(xn.shape) < 0.4] = 0.0
        xd = torch.from_numpy(xn).cuda()
        wd = torch.from_numpy(grid_values(E, M, bW, (Co, C // groups, k, k), 13, 0.8)).cuda()
        xb, rb = torch.tensor([float(bA)], device="cuda"), torch.tensor([8.0], device="cuda")
        wb = torch.full((Co,), float(bW), device="cuda")
        call = lambda: fq.ops.approx_conv2d(xd, wd, xb, wb, rb, fq.ops.make_params(MAIN), (st, st), (pad, pad), (dil, dil), groups)
        os.environ["B200_LUT_SPLITK"] = "1"
        os.environ["B200_CONV_GATHER"] = "2"
        try:
            want = call()
            got, path = _streamk(call)
        finally:
            os.environ.pop("B200_LUT_SPLITK", None)
            os.environ.pop("B200_CONV_GATHER", None)
        assert path == "lut-streamk", (path, case)
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), case
