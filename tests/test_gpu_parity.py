"""GPU parity tests (run on a B200 with `-m gpu`): the CUDA path, called through the C ABI
(ctypes -> libfp8approx_b200.so), against (a) the golden vectors produced by the unmodified reference
and (b) the CPU oracle on the same seeded inputs.  Integer fields and K=1 addends: bit exact.
GEMMs: bit exact against the oracle (same k-ascending fp32 order) and against the reference
whenever fp32 summation is exact, else norm-wise 1e-6 (the tolerance north_star states)."""
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
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b)) | ((a == 0) & (b == 0))))


def signed_codes(E, M, bias, extra_top=True):
    n = 2 ** (E + M) + (2 ** M if extra_top else 0)
    c = np.arange(n, dtype=np.int32)
    v = orc.decode(c >> M, c & (2 ** M - 1), E, M, bias)
    return np.concatenate([v, -v]).astype(np.float32)


def params_from_row(fq, row, **kw):
    E, M, wc, dn, bA, bB, bR, wa, s2, qb, clip, tb = [int(v) for v in row]
    p = fq.ops.make_params(dict(expo_width=E, mant_width=M, withComp=bool(wc), dnsmp_factor=dn, with_approx=bool(wa),
                                with_s2nn2s_opt=bool(s2), quant_btw_mult_accu=bool(qb), golden_clip_OF=bool(clip)),
                           single_column_quirk=bool(tb), **kw)
    okw = dict(E=E, M=M, with_approx=wa, with_s2nn2s_opt=s2, quant_btw_mult_accu=qb, golden_clip_OF=clip,
               tensor_bias_quirk=tb)
    return p, okw, orc.get_error_table_NN(E, M, bool(wc), dn), bA, bB, bR


@pytest.mark.parametrize("fmt", [(4, 3), (3, 4), (2, 5)])
@pytest.mark.parametrize("force_generic", [False, True])
def test_exhaustive_pairs(fq, fmt, force_generic):
    """All signed codes x all signed codes, K = 1, every fixture flag/bias combination: bit exact
    against the reference's own outputs, for the table kernel and the arithmetic kernel."""
    E, M = fmt
    z = np.load(os.path.join(GOLD, f"pairs_E{E}M{M}.npz"))
    bad = []
    paths = set()
    for i, row in enumerate(z["cfgs"]):
        p, okw, T, bA, bB, bR = params_from_row(fq, row, force_generic=force_generic)
        a = dev(signed_codes(E, M, bA).reshape(-1, 1))
        b = dev(signed_codes(E, M, bB).reshape(1, -1))
        got = fq.ops.approx_matmul(a, b, bA, bB, bR, p).cpu().numpy()
        paths.add(fq.ops.last_path())
        if not same_bits(got, z[f"out{i}"]):
            d = np.argwhere(got.view(np.uint32) != z[f"out{i}"].view(np.uint32))
            bad.append((i, row.tolist(), len(d), d[:3].tolist()))
    assert not bad, bad[:5]
    if not force_generic:
        assert "lut" in paths     # the table kernel really ran for the main mode


def test_debug_params_fixture(fq):
    """The reference's own dumped MobileNetV2 layer (debug_params/*.csv), 32 columns with per-column
    weight bias, against the reference's column-loop output."""
    z = np.load(os.path.join(GOLD, "debug_params.npz"))
    A, B, bB, bA = z["act"], z["weight"], z["weight_bias"], int(z["act_bias"])
    for i, row in enumerate(z["cfgs"]):
        p, okw, T, _, _, bR = params_from_row(fq, row)
        got = fq.ops.approx_matmul(dev(A), dev(B), bA, dev(bB.astype(np.int32)), bR, p).cpu().numpy()
        want = orc.approx_matmul(A, B, T, bA, bB, bR, **okw)
        assert same_bits(got, want), ("oracle", row.tolist())
        if okw["quant_btw_mult_accu"]:
            assert same_bits(got, z[f"out{i}"]), ("reference", row.tolist())
        else:
            ab = orc.approx_matmul_abs(A, B, T, bA, bB, bR, **okw)
            assert np.all(np.abs(got.astype(np.float64) - z[f"out{i}"]) <= 1e-6 * ab + 1e-30)


def test_gemm_random_fixture(fq):
    z = np.load(os.path.join(GOLD, "gemm_random.npz"))
    for i, row in enumerate(z["cfgs"]):
        p, okw, T, bA, _, bR = params_from_row(fq, row)
        A, B, bB = z[f"A{i}"], z[f"B{i}"], z[f"bB{i}"]
        got = fq.ops.approx_matmul(dev(A), dev(B), bA, dev(bB), bR, p).cpu().numpy()
        want = orc.approx_matmul(A, B, T, bA, bB, bR, **okw)
        assert same_bits(got, want), row.tolist()
        ab = orc.approx_matmul_abs(A, B, T, bA, bB, bR, **okw)
        assert np.all(np.abs(got.astype(np.float64) - z[f"out{i}"]) <= 1e-6 * ab + 1e-30)


def _grid_sample(rng, E, M, bias, shape, zero_frac, spread=None):
    vals = signed_codes(E, M, bias, extra_top=False)
    if spread is not None:   # keep the top `spread` binades only (realistic weights)
        vals = vals[np.abs(vals) >= np.abs(vals).max() / 2 ** spread]
    x = rng.choice(vals, size=shape).astype(np.float32)
    x[rng.random(shape) < zero_frac] = 0.0
    return x


# (E, M, dnsmp_factor, expected kernel path).  E2M5 / withComp / dnsmp 3 selects the one table whose
# pair function is not odd in the operand signs (4 cells at the round-to-zero edge, v9 :68 after :36):
# the library detects that when it builds the pair table and switches to the signed-index table
# layout (both operand signs in the table index, csrc/lut.cuh).
@pytest.mark.parametrize("fmt", [(4, 3, 3, "lut"), (3, 4, 3, "lut"), (2, 5, 4, "lut"), (2, 5, 3, "lut-signed")])
@pytest.mark.parametrize("shape", [(300, 200, 130), (129, 1000, 257), (64, 4096, 40)])
def test_gemm_vs_oracle(fq, fmt, shape):
    """Ragged shapes (not multiples of the tile), weight.t() views, per-column bias: table kernel vs
    oracle, bit exact; and the arithmetic kernel must agree with the table kernel."""
    E, M, dnsmp, path = fmt
    Mr, K, N = shape
    rng = np.random.default_rng(1000 * E + K)
    half = 2 ** (E - 1)
    bA, bR = half + 1, half - 1
    bB = rng.integers(half, half + 4, size=N).astype(np.int32)
    A = _grid_sample(rng, E, M, bA, (Mr, K), 0.3)
    W = np.stack([_grid_sample(rng, E, M, int(bB[j]), (K,), 0.02) for j in range(N)], axis=0)   # [N,K] weights
    T = orc.get_error_table_NN(E, M, True, dnsmp)
    p = fq.ops.make_params(dict(expo_width=E, mant_width=M, withComp=True, dnsmp_factor=dnsmp, with_approx=True,
                                with_s2nn2s_opt=True, quant_btw_mult_accu=True))
    got = fq.ops.approx_matmul(dev(A), dev(W).t(), bA, dev(bB), bR, p).cpu().numpy()
    assert fq.ops.last_path() in (path, path + "-splitk", path + "-streamk")
    want = orc.approx_matmul(A, W.T.copy(), T, bA, bB, bR, E=E, M=M, with_approx=1, with_s2nn2s_opt=1,
                             quant_btw_mult_accu=1)
    assert same_bits(got, want)
    pg = fq.ops.make_params(dict(expo_width=E, mant_width=M, withComp=True, dnsmp_factor=dnsmp, with_approx=True,
                                 with_s2nn2s_opt=True, quant_btw_mult_accu=True), force_generic=True)
    got2 = fq.ops.approx_matmul(dev(A), dev(W).t(), bA, dev(bB), bR, pg).cpu().numpy()
    assert fq.ops.last_path() == "generic"
    assert same_bits(got2, want)


def test_offgrid_falls_back_on_device(fq):
    """One off-grid operand anywhere: the device-side flag routes the call to the arithmetic kernel;
    the result is still the reference's (oracle) answer."""
    E, M = 4, 3
    rng = np.random.default_rng(7)
    A = _grid_sample(rng, E, M, 9, (70, 96), 0.2)
    B = _grid_sample(rng, E, M, 10, (96, 33), 0.0)
    A[13, 57] = np.float32(0.3337)     # not representable with 3 mantissa bits
    bB = np.full(33, 10, np.int32)
    T = orc.get_error_table_NN(E, M, False, 3)
    p = fq.ops.make_params(dict(expo_width=E, mant_width=M, withComp=False, with_approx=True, with_s2nn2s_opt=True,
                                quant_btw_mult_accu=True))
    got = fq.ops.approx_matmul(dev(A), dev(B), 9, dev(bB), 8, p).cpu().numpy()
    want = orc.approx_matmul(A, B, T, 9, bB, 8, E=E, M=M, with_approx=1, with_s2nn2s_opt=1, quant_btw_mult_accu=1)
    assert same_bits(got, want)


def test_empty_and_errors(fq):
    p = fq.ops.make_params(dict(expo_width=4, mant_width=3))
    out = fq.ops.approx_matmul(torch.zeros(0, 8).cuda(), torch.zeros(8, 4).cuda(), 8, 8, 8, p)
    assert out.shape == (0, 4)
    out = fq.ops.approx_matmul(torch.zeros(5, 0).cuda(), torch.zeros(0, 4).cuda(), 8, 8, 8, p)
    assert out.shape == (5, 4) and float(out.abs().sum()) == 0.0
    with pytest.raises(ValueError):      # approx_matmul_whole_v9.py:590
        fq.ops.approx_matmul(torch.zeros(2, 2).cuda(), torch.zeros(2, 2).cuda(), 8, 8, 8,
                             fq.ops.make_params(dict(expo_width=5, mant_width=2)))
    with pytest.raises(UnboundLocalError):   # :566-572
        fq.ops.approx_matmul(torch.zeros(2, 2).cuda(), torch.zeros(2, 2).cuda(), 8, 8, 8,
                             fq.ops.make_params(dict(expo_width=3, mant_width=4, withComp=True, dnsmp_factor=2)))
    with pytest.raises(AssertionError):      # :20
        fq.custom_matmul_vectorize(torch.zeros(2, 3).cuda(), torch.zeros(2, 2).cuda(), 4, 3, 8, 8, 8, None)


def test_custom_matmul_vectorize_mirror(fq):
    """Reference-shaped call (table tensor argument, shared biases, hand tensors of
    approx_matmul_whole_v9.py:836-846)."""
    E, M = 3, 4
    Iact = np.array([[0.5625, 0.3438, 0.1328125, 0.0859375], [0.0625, 0.3750, 0.4375, 0.5],
                     [0.96875, 1, 1.0625, 1.125]], np.float32)
    Wght = np.array([[-0.109375], [-0.2421875], [0.05859375], [0.2109375]], np.float32)
    Iq = orc.quant(Iact, E, M, 3, clip_OF=True)
    Wq = orc.quant(Wght, E, M, 5, clip_OF=True)
    for (wc, dn) in [(True, 3), (True, 5), (False, 3)]:
        T = orc.get_error_table_NN(E, M, wc, dn)
        table = fq.get_error_table_NN(E, M, wc, dn)
        assert np.array_equal(table.cpu().numpy(), T)
        for s2 in (False, True):
            got = fq.custom_matmul_vectorize(dev(Iq), dev(Wq), E, M, 3, 5, 5, table, with_approx=True,
                                             with_s2nn2s_opt=s2, quant_btw_mult_accu=True).cpu().numpy()
            want = orc.approx_matmul(Iq, Wq, T, 3, 5, 5, E=E, M=M, with_approx=1, with_s2nn2s_opt=int(s2),
                                     quant_btw_mult_accu=1)
            assert same_bits(got, want)


def test_fakequant_fixture(fq):
    z = np.load(os.path.join(GOLD, "fakequant.npz"))
    for i, (mbits, sbits, pc) in enumerate(z["cfgs"]):
        y, b = fq.quantize_to_fp8_ste_MM(dev(z[f"x{i}"]), 8, dev(z[f"mv{i}"]), torch.tensor([float(mbits)]), int(sbits))
        assert np.array_equal(b.cpu().numpy().reshape(-1), z[f"b{i}"].reshape(-1)), i
        assert tuple(b.shape) == tuple(z[f"b{i}"].shape), (i, b.shape, z[f"b{i}"].shape)
        assert same_bits(y.cpu().numpy(), z[f"y{i}"]), i


def test_fakequant_special_values(fq):
    """NaN propagates (torch.min / torch.max do), +-Inf clamps to +-maxval, fp32 sub-normals, signed zeros: outputs of
    the unmodified reference (tests/golden/make_fakequant_special.py), stand-alone kernel and fused GEMM prologue."""
    z = np.load(os.path.join(GOLD, "fakequant_special.npz"))
    for i, (mbits, sbits) in enumerate(z["cfgs"]):
        y, b = fq.quantize_to_fp8_ste_MM(dev(z["x"]), 8, dev(z[f"mv{i}"]), torch.tensor([float(mbits)]), int(sbits))
        assert np.array_equal(b.cpu().numpy().reshape(-1), z[f"b{i}"].reshape(-1)), i
        got, want = y.cpu().numpy(), z[f"y{i}"]
        ok = (got.view(np.uint32) == want.view(np.uint32)) | ((got == 0) & (want == 0)) | (np.isnan(got) & np.isnan(want))
        assert ok.all(), (i, got, want)


@pytest.mark.parametrize("mbits", [2, 3, 4, 5])
def test_fakequant_large_vs_oracle(fq, mbits):
    """16 M elements (> L2 is not needed for parity): per tensor and per channel, values straddling
    powers of two by +-1 ulp included (the log2 rounding hazard of fp8_quantizer.py:144)."""
    rng = np.random.default_rng(mbits)
    x = (rng.standard_normal((64, 1 << 14)) * np.exp(rng.standard_normal((64, 1)) * 2)).astype(np.float32)
    p2 = (np.float32(2.0) ** rng.integers(-14, 6, size=4096)).astype(np.float32)
    x[0, :4096] = np.nextafter(p2, np.float32(0))
    x[1, :4096] = np.nextafter(p2, np.float32(100))
    x[2, :4096] = p2
    mv = np.array([np.abs(x).max()], np.float32)
    y, b = fq.quantize_to_fp8_ste_MM(dev(x), 8, dev(mv), torch.tensor([float(mbits)]), 1)
    yo, bo = orc.fp8_fakequant(x, mv, 8, float(mbits), 1)
    assert np.array_equal(b.cpu().numpy().reshape(-1), bo) and same_bits(y.cpu().numpy(), yo)
    mvc = np.abs(x).max(axis=1).astype(np.float32)
    y, b = fq.quantize_to_fp8_ste_MM(dev(x), 8, dev(mvc), torch.tensor([float(mbits)]), 1)
    yo, bo = orc.fp8_fakequant(x, mvc, 8, float(mbits), 1)
    assert np.array_equal(b.cpu().numpy().reshape(-1), bo) and same_bits(y.cpu().numpy(), yo)
    # idempotence (size-independent property): quantizing a quantized tensor changes nothing
    y2, _ = fq.quantize_to_fp8_ste_MM(y, 8, dev(mvc), torch.tensor([float(mbits)]), 1)
    assert torch.equal(y2, y)


def test_codec_fixture(fq):
    z = np.load(os.path.join(GOLD, "codec.npz"))
    for i, (E, M, bias, clip) in enumerate(z["cfgs"]):
        E, M, bias, clip = int(E), int(M), int(bias), bool(clip)
        x = dev(z[f"x{i}"])
        e, m = fq.ops.fpany_encode(x, E, M, bias, clip_OF=clip)
        assert np.array_equal(e.cpu().numpy(), z[f"e{i}"]) and np.array_equal(m.cpu().numpy(), z[f"m{i}"])
        q = fq.quant_to_fp_any_vectorize_torch(x, E, M, custom_bias=bias, clip_OF=clip)
        assert same_bits(q.cpu().numpy(), z[f"q{i}"])
        e5, m5 = fq.ops.fpany_encode(x, E, M, bias, clip_OF=clip, allnorm=True)
        assert np.array_equal(e5.cpu().numpy(), z[f"e5_{i}"]) and np.array_equal(m5.cpu().numpy(), z[f"m5_{i}"])
        d5 = fq.ops.fpany_decode(e5, m5, E, M, bias, allnorm=True)
        assert same_bits(d5.cpu().numpy(), z[f"d5_{i}"])
        # decode(encode(v)) == quant(v) for the sub-normal aware codec
        d = fq.ops.fpany_decode(e, m, E, M, bias, sign=torch.where(x < 0, -1.0, 1.0))
        assert same_bits(d.cpu().numpy(), z[f"q{i}"])


def test_v5_ofuf_fixture(fq):
    """Row A7 on the GPU: b200_approx_mult_v5 against the unmodified reference's approx_mult_new
    (integer codes: bit exact), with explicit mantissa tensors and with them derived on the device."""
    z = np.load(os.path.join(GOLD, "v5_ofuf.npz"))
    for i, (E, M, wc, dn, B_neg, sim, of, uf, g) in enumerate(z["cfgs"]):
        E, M = int(E), int(M)
        x, y, T = dev(z[f"x{g}"]), dev(z[f"y{g}"]), dev(z[f"t{g}"])
        got = fq.ops.approx_mult_v5(x, y, int(B_neg), x % (2 ** M), y % (2 ** M), T, sim, of, uf, E, M)
        assert np.array_equal(got.cpu().numpy(), z[f"r{i}"]), (i, E, M, sim, of, uf)
        if i % 8 == 7:
            got = fq.ops.approx_mult_v5(x, y, int(B_neg), None, None, T, sim, of, uf, E, M)
            assert np.array_equal(got.cpu().numpy(), z[f"r{i}"])


@pytest.fixture(params=["1", "0"], ids=["decoded-plane", "per-mac-decode"])
def dw_kernel(request):
    """Both single-column table kernels: the decoded-plane one (default) and the older per-MAC decoding one."""
    os.environ["B200_DW_PLANE"] = request.param
    yield "single-column-lut-plane" if request.param == "1" else "single-column-lut"
    os.environ.pop("B200_DW_PLANE", None)


@pytest.mark.parametrize("fmt", [(4, 3, 3), (3, 4, 3), (2, 5, 3), (2, 5, 5)])
@pytest.mark.parametrize("geom", [(3, 40, 17, 19, 3, 1, 1), (2, 33, 16, 16, 3, 2, 1), (1, 8, 9, 9, 5, 1, 2)])
def test_depthwise_table_kernel_vs_arithmetic(fq, fmt, geom, dw_kernel):
    """Single-column groups (every depthwise conv): the table kernel against the arithmetic kernel that the
    layer fixtures pin to the reference's tensor-bias branch (approx_calculation.py:800-809): bit exact.
    Also: biases < 2 and off-grid operands must fall back on the device."""
    E, M, dnsmp = fmt
    B, C, H, W, k, stride, pad = geom
    rng = np.random.default_rng(100 * E + C)
    half = 2 ** (E - 1)
    bA, bR = half + 1, half + 2
    bW = rng.integers(max(2, half), half + 4, size=C).astype(np.float32)
    x = _grid_sample(rng, E, M, bA, (B, C, H, W), 0.4)
    w = np.stack([_grid_sample(rng, E, M, int(bW[c]), (1, k, k), 0.05) for c in range(C)], axis=0)
    kw = dict(expo_width=E, mant_width=M, withComp=True, dnsmp_factor=dnsmp, with_approx=True, with_s2nn2s_opt=True,
              quant_btw_mult_accu=True)
    args = (dev(x), dev(w), torch.tensor([float(bA)]).cuda(), dev(bW), torch.tensor([float(bR)]).cuda())
    conv = ((stride, stride), (pad, pad), (1, 1), C)
    got = fq.ops.approx_conv2d(*args, fq.ops.make_params(kw, single_column_quirk=True), *conv)
    assert fq.ops.last_path() == dw_kernel
    want = fq.ops.approx_conv2d(*args, fq.ops.make_params(kw, single_column_quirk=True, force_generic=True), *conv)
    assert fq.ops.last_path() == "single-column"
    assert same_bits(got.cpu().numpy(), want.cpu().numpy())
    assert float(want.abs().max()) > 0
    # a result bias of 1 leaves the table's domain (integer pow gives min_norm = 1, not 0)
    args1 = args[:4] + (torch.tensor([1.0]).cuda(),)
    got = fq.ops.approx_conv2d(*args1, fq.ops.make_params(kw, single_column_quirk=True), *conv)
    want = fq.ops.approx_conv2d(*args1, fq.ops.make_params(kw, single_column_quirk=True, force_generic=True), *conv)
    assert same_bits(got.cpu().numpy(), want.cpu().numpy())
    # an off-grid activation
    x2 = x.copy()
    x2[0, 0, 1, 1] = np.float32(0.3337)
    args2 = (dev(x2),) + args[1:]
    got = fq.ops.approx_conv2d(*args2, fq.ops.make_params(kw, single_column_quirk=True), *conv)
    want = fq.ops.approx_conv2d(*args2, fq.ops.make_params(kw, single_column_quirk=True, force_generic=True), *conv)
    assert same_bits(got.cpu().numpy(), want.cpu().numpy())
