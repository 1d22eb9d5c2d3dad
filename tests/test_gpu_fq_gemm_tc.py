"""GPU tests of the non-approx matmul on the tensor cores (K3, csrc/fq_gemm_tc.cu), through the C ABI.

The reference computes `x @ y[:, i]` per column in fp32 (approx/approx_calculation.py:968-969); the
kernel multiplies exact bf16 copies of the fake-quantized operands and accumulates in fp32 in tensor
memory.  Products are exact, so the only difference is the summation order.  Checks:
  * exact equality with an fp64 reference wherever the fp32 sum is exact (small-integer grids);
  * the tolerance north_star states elsewhere: |C - ref| <= 1e-6 * sum_k |a_k b_k|  (norm-wise);
  * ragged shapes (TMA zero fill of out-of-range rows / k), strided right operands, bias epilogue;
  * operands that are NOT exact in bf16 fall back ON THE DEVICE to the fp32 CUDA-core kernel.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fq():
    import fp8_quantization_b200 as pkg
    pkg._lib.load()
    assert torch.cuda.is_available()
    return pkg


def _fakequant(fq, t, M, per_channel):
    mv = t.abs().amax(dim=1) if per_channel else t.abs().max().reshape(1)
    return fq.ops.fp8_fakequant(t, 8, mv, torch.tensor([float(M)]), 1)[0]


def _check(got, a, b, bias=None, tol=1e-6):
    a64, b64 = a.double(), b.double()
    ref = a64 @ b64
    bound = a64.abs() @ b64.abs()
    if bias is not None:
        ref = ref + bias.double()
        bound = bound + bias.double().abs()
    err = (got.double() - ref).abs()
    assert bool((err <= tol * bound + 1e-30).all()), float((err / (bound + 1e-30)).max())


@pytest.mark.parametrize("shape", [(128, 64, 128), (256, 512, 256), (300, 200, 130), (129, 1000, 257), (64, 4096, 40),
                                   (1, 8, 1), (5, 1, 3), (1000, 77, 513)])
@pytest.mark.parametrize("M", [3, 5])
def test_fq_gemm_tc_vs_fp64(fq, shape, M):
    Mr, K, N = shape
    g = torch.Generator().manual_seed(100 * K + N)
    x = _fakequant(fq, (torch.randn(Mr, K, generator=g) * 0.5).cuda(), M, False)
    w = _fakequant(fq, (torch.randn(N, K, generator=g) * 0.05).cuda(), M, True)
    bias = (torch.randn(N, generator=g) * 0.01).cuda()
    got = fq.ops.fq_matmul(x, w.t())                       # weight.t(): k contiguous
    assert fq.ops.last_path() == "tcgen05"
    _check(got, x, w.t())
    got_b = fq.ops.fq_matmul(x, w.t().contiguous(), bias=bias)   # [K, N] row-major right operand + bias epilogue
    _check(got_b, x, w.t(), bias)


def test_fq_gemm_tc_exact_on_integers(fq):
    """Small integers: every partial sum is exact in fp32, so any summation order gives the same
    bits -- the tensor-core result must EQUAL the exact product (checks descriptors, swizzle, the
    TMEM lane/column mapping and the k-advance inside the swizzle row)."""
    g = torch.Generator().manual_seed(3)
    for (Mr, K, N) in [(128, 64, 256), (384, 320, 512), (130, 100, 300)]:
        a = torch.randint(-8, 9, (Mr, K), generator=g).float().cuda()
        b = torch.randint(-8, 9, (K, N), generator=g).float().cuda()
        got = fq.ops.fq_matmul(a, b)
        want = (a.double() @ b.double()).float()
        assert torch.equal(got, want), (Mr, K, N, float((got - want).abs().max()))


def test_bias_view_that_is_not_16_byte_aligned(fq):
    """A sliced bias (data_ptr 4 bytes off a 16-byte boundary) must not be read with 128-bit loads: the epilogue falls
    back to scalar bias loads (the TMA-store and the vector-store paths both require an aligned bias)."""
    g = torch.Generator().manual_seed(4)
    for (Mr, K, N) in [(128, 64, 256), (300, 96, 512)]:
        a = torch.randint(-8, 9, (Mr, K), generator=g).float().cuda()
        b = torch.randint(-8, 9, (K, N), generator=g).float().cuda()
        store = torch.randint(-8, 9, (N + 4,), generator=g).float().cuda()
        bias = store[1:N + 1]
        assert bias.data_ptr() % 16 == 4
        got = fq.ops.fq_matmul(a, b, bias=bias)
        want = (a.double() @ b.double() + bias.double()).float()
        assert torch.equal(got, want)


@pytest.mark.parametrize("bn", ["128", "192", "256"])
def test_tile_widths_return_the_same_bits(fq, bn, monkeypatch):
    """128 x {128, 192, 256} tiles of the single-CTA kernel (chosen by SM fill; B200_TC_BN forces one): same k order per
    output element, hence the same bits -- with a bias, whose registers are laid out per tile width."""
    g = torch.Generator().manual_seed(9)
    monkeypatch.setenv("B200_TC_PAIR", "0")
    for (Mr, K, N) in [(300, 96, 520), (1024, 128, 768), (130, 64, 200)]:
        a = torch.randint(-8, 9, (Mr, K), generator=g).float().cuda()
        b = torch.randint(-8, 9, (K, N), generator=g).float().cuda()
        bias = torch.randint(-8, 9, (N,), generator=g).float().cuda()
        want = (a.double() @ b.double() + bias.double()).float()
        monkeypatch.setenv("B200_TC_BN", bn)
        got = fq.ops.fq_matmul(a, b, bias=bias)
        assert fq.ops.last_path() == "tcgen05"
        assert torch.equal(got, want), (Mr, K, N, bn)


def test_fq_gemm_tc_offgrid_falls_back_on_device(fq):
    g = torch.Generator().manual_seed(4)
    a = torch.randn(70, 96, generator=g).cuda()            # full fp32 mantissas: not exact in bf16
    b = torch.randn(96, 33, generator=g).cuda()
    got = fq.ops.fq_matmul(a, b)
    _check(got, a, b)
    ref = a @ b
    assert float((got - ref).abs().max()) <= 1e-5 * float(ref.abs().max())


def test_fq_gemm_tc_empty(fq):
    out = fq.ops.fq_matmul(torch.zeros(0, 8).cuda(), torch.zeros(8, 4).cuda())
    assert out.shape == (0, 4)
    bias = torch.arange(4.0).cuda()
    out = fq.ops.fq_matmul(torch.zeros(5, 0).cuda(), torch.zeros(0, 4).cuda(), bias=bias)
    assert out.shape == (5, 4) and torch.equal(out, bias.expand(5, 4))


def test_plain_linear_layer_uses_tensor_cores(fq):
    """QCustomLinearTorch with approx_flag=False (plain `A @ B`, the res_quantizer
    skipped) with the matmul on tcgen05; compared with the same layer math in fp64 torch."""
    from fp8_quantization_b200.approx_calculation import QCustomLinearTorch
    torch.manual_seed(10)
    layer = QCustomLinearTorch(
        192, 160, bias=True, method="fp_quantizer", act_method="fp_quantizer", weight_range_method="current_minmax",
        act_range_method="allminmax", n_bits=8, n_bits_act=8, per_channel_weights=True, quantize_input=True,
        fp8_kwargs=dict(mantissa_bits=3, set_maxval=True),
        custom_approx_params=dict(expo_width=4, mant_width=3),
        run_method=dict(approx_flag=False, quantize_after_mult_and_add=False, res_quantizer_flag=False,
                        original_quantize_res=True)).cuda().eval()
    layer.quantized()
    x = torch.randn(3, 50, 192).cuda()                     # 3-D input: flattened to rows
    with torch.inference_mode():
        layer.estimate_ranges()
        layer(x)
        layer.fix_ranges()
        y = layer(x)
        assert fq.ops.last_path() == "tcgen05"
        xq = layer.activation_quantizer(x)
        wq, b = layer.get_params()
        ref = (xq.double().reshape(-1, 192) @ wq.double().t() + b.double()).reshape(3, 50, 160)
    assert float((y.double() - ref).abs().max()) <= 1e-5 * float(ref.abs().max())


@pytest.mark.parametrize("shape", [(300, 200, 130), (129, 1000, 257), (512, 768, 768)])
def test_prepared_right_operand_equals_per_call_conversion(fq, shape):
    """b200_fq_gemm_tc_prepare_b + b200_fq_gemm_tc_pre (opt-in: the bf16 plane of fixed-range weights built once) return
    the bits of b200_fq_gemm_tc, for both operand layouts, with the bias epilogue, on repeated calls; a right operand
    that is not exact in bf16 carries its flag in the plane header and still takes the fp32 fallback."""
    Mr, K, N = shape
    g = torch.Generator().manual_seed(7 * K + N)
    x = _fakequant(fq, (torch.randn(Mr, K, generator=g) * 0.5).cuda(), 3, False)
    w = _fakequant(fq, (torch.randn(N, K, generator=g) * 0.05).cuda(), 3, True)
    bias = (torch.randn(N, generator=g) * 0.01).cuda()
    for y in (w.t(), w.t().contiguous()):
        want = fq.ops.fq_matmul(x, y, bias=bias)
        assert fq.ops.last_path() == "tcgen05"
        for _ in range(2):
            got = fq.ops.fq_matmul(x, y, bias=bias, cache_b=True)
            assert fq.ops.last_path() == "tcgen05-preB"
            assert torch.equal(got.view(torch.int32), want.view(torch.int32))
        got2 = fq.ops.fq_matmul(x * 0.5, y, cache_b=True)           # same plane, other activations
        assert torch.equal(got2.view(torch.int32), fq.ops.fq_matmul(x * 0.5, y).view(torch.int32))
    w_bad = w.clone()
    w_bad[3, 5] = 0.1234567                                          # not exact in bf16
    got = fq.ops.fq_matmul(x, w_bad.t(), cache_b=True)
    want = fq.ops.fq_matmul(x, w_bad.t())
    assert torch.equal(got.view(torch.int32), want.view(torch.int32))
    _check(got, x, w_bad.t())


@pytest.mark.parametrize("shape", [(300, 200, 130), (129, 1000, 257), (512, 768, 768), (1000, 77, 513), (2048, 4096, 1024)])
def test_cta_pair_kernel_equals_single_cta_kernel(fq, shape):
    """fq_gemm_tc2_kernel (cta_group::2: a CTA pair per 256 x 256 tile, B shared across the two SMs) against the
    single-CTA kernel: same products, same fp32 accumulation in tensor memory over the same k order -> same bits;
    ragged rows / columns / K through TMA zero fill, bias epilogue, prepared right operand."""
    import os
    Mr, K, N = shape
    g = torch.Generator().manual_seed(3 * K + N)
    x = _fakequant(fq, (torch.randn(Mr, K, generator=g) * 0.5).cuda(), 3, False)
    w = _fakequant(fq, (torch.randn(N, K, generator=g) * 0.05).cuda(), 3, True)
    bias = (torch.randn(N, generator=g) * 0.01).cuda()
    outs = {}
    for mode in ("0", "2"):
        os.environ["B200_TC_PAIR"] = mode
        try:
            outs[mode] = (fq.ops.fq_matmul(x, w.t(), bias=bias), fq.ops.last_path(),
                          fq.ops.fq_matmul(x, w.t(), cache_b=True), fq.ops.last_path())
        finally:
            os.environ.pop("B200_TC_PAIR", None)
    assert outs["0"][1] == "tcgen05" and outs["2"][1] == "tcgen05-pair", (outs["0"][1], outs["2"][1])
    assert outs["2"][3] == "tcgen05-pair-preB"
    _check(outs["2"][0], x, w.t(), bias)
    assert torch.equal(outs["0"][0].view(torch.int32), outs["2"][0].view(torch.int32))
    assert torch.equal(outs["0"][2].view(torch.int32), outs["2"][2].view(torch.int32))
