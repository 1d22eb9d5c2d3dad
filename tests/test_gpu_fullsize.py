"""Full-size parity (BASELINE.json configs[1]: 4096 rows x K 4096 x N 4096, every shipped format, both
modes).  The oracle cannot finish 68.7 G pair operations in seconds, so the full launch is checked
through size-independent properties of the path, each bit exact:

  * a random row sample of the full launch against the CPU oracle (all 4096 columns of those rows);
  * the table kernel against the arithmetic kernel (`pair_generic`, the definition the exhaustive
    code x code fixtures pin to the reference) over the WHOLE output;
  * row independence: permuting activation rows permutes output rows (approx_matmul_whole_v9.py has no
    cross-row term; this is what the multi-GPU row sharding relies on);
  * column independence: permuting weight rows together with their per-channel biases permutes columns;
  * zero padding along K: appended zero activations contribute nothing (approx_matmul_whole_v9.py:81),
    whatever the weights there are;
  * K-split additivity wherever fp32 summation is exact (every addend is a multiple of the result
    sub-normal spacing u and |sum|/u < 2^24): C = C[:, :K/2] + C[:, K/2:].
"""
import numpy as np
import pytest
import torch

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

ROWS = K_DIM = N_DIM = 4096
SEED = 10


@pytest.fixture(scope="module")
def fq():
    import fp8_quantization_b200 as pkg
    pkg._lib.load()
    assert torch.cuda.is_available()
    return pkg


def same_bits(a, b):
    if torch.is_tensor(a):
        return bool(torch.all((a.view(torch.int32) == b.view(torch.int32)) | ((a == 0) & (b == 0))))
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0))))


def operands(fq, M, zero_frac=0.0):
    """SURVEY.md 8(d) synthetic operands, fake-quantized by the path's own quantizer kernel."""
    g = torch.Generator().manual_seed(SEED + M)
    x = (torch.randn(ROWS, K_DIM, generator=g) * 0.5)
    w = (torch.randn(N_DIM, K_DIM, generator=g) * 0.02)
    if zero_frac:
        x[torch.rand(x.shape, generator=g) < zero_frac] = 0.0
    x, w = x.cuda(), w.cuda()
    mant = torch.tensor([float(M)])
    xq, bA = fq.quantize_to_fp8_ste_MM(x, 8, x.abs().max().reshape(1), mant, 1)
    wq, bB = fq.quantize_to_fp8_ste_MM(w, 8, w.abs().amax(dim=1), mant, 1)
    exact_max = float((xq[:256] @ wq.t()).abs().max()) * 1.5
    return xq, wq, bA, bB.reshape(-1), exact_max


def approx_kw(M):
    return dict(expo_width=7 - M, mant_width=M, withComp=True, dnsmp_factor=3, with_approx=True,
                with_s2nn2s_opt=True, quant_btw_mult_accu=True)


@pytest.mark.parametrize("M", [3, 4, 5])
def test_fullsize_approx_properties(fq, M):
    E = 7 - M
    xq, wq, bA, bB, exact_max = operands(fq, M, zero_frac=0.25 if M == 4 else 0.0)
    bR = torch.tensor([float(orc.fp8_bias(exact_max, 8, float(M), 1))]).cuda()
    p = fq.ops.make_params(approx_kw(M))
    C = fq.ops.approx_matmul(xq, wq.t(), bA, bB, bR, p)
    path = fq.ops.last_path()
    assert path.startswith("lut"), path
    assert float(C.abs().max()) > 0

    # (1) oracle on a row sample, every column
    rng = np.random.default_rng(M)
    rows = np.sort(rng.choice(ROWS, size=12, replace=False))
    T = orc.get_error_table_NN(E, M, True, 3)
    want = orc.approx_matmul(xq[rows].cpu().numpy(), np.ascontiguousarray(wq.t().cpu().numpy()), T, int(bA.item()),
                             bB.to(torch.int32).cpu().numpy(), int(bR.item()), E=E, M=M, with_approx=1,
                             with_s2nn2s_opt=1, quant_btw_mult_accu=1)
    assert same_bits(C[rows].cpu().numpy(), want)

    # (2) table kernel == arithmetic kernel over the whole output
    Cg = fq.ops.approx_matmul(xq, wq.t(), bA, bB, bR, fq.ops.make_params(approx_kw(M), force_generic=True))
    assert fq.ops.last_path() == "generic"
    assert same_bits(C, Cg)
    del Cg

    # (3) row permutation
    g = torch.Generator(device="cuda").manual_seed(1)
    pr = torch.randperm(ROWS, device="cuda", generator=g)
    assert same_bits(fq.ops.approx_matmul(xq[pr].contiguous(), wq.t(), bA, bB, bR, p), C[pr])
    # (4) column permutation (weights and their per-channel exponent biases move together)
    pc = torch.randperm(N_DIM, device="cuda", generator=g)
    assert same_bits(fq.ops.approx_matmul(xq, wq[pc].contiguous().t(), bA, bB[pc].contiguous(), bR, p), C[:, pc])
    # (5) zero padding along K, arbitrary (on-grid) weights under the zeros
    pad = 160
    xz = torch.cat([xq, torch.zeros(ROWS, pad, device="cuda")], dim=1).contiguous()
    wz = torch.cat([wq, wq[:, :pad]], dim=1).contiguous()
    assert same_bits(fq.ops.approx_matmul(xz, wz.t(), bA, bB, bR, p), C)
    del xz, wz
    # (6) K-split additivity where fp32 summation is exact
    u = 2.0 ** (1 - int(bR.item()) - M)
    h = K_DIM // 2
    C1 = fq.ops.approx_matmul(xq[:, :h].contiguous(), wq[:, :h].contiguous().t(), bA, bB, bR, p)
    C2 = fq.ops.approx_matmul(xq[:, h:].contiguous(), wq[:, h:].contiguous().t(), bA, bB, bR, p)
    for t in (C, C1, C2):     # every partial sum is an integer multiple of u
        q = t.double() / u
        assert bool(torch.all(q == q.round()))
    bound = float(torch.maximum(C1.abs(), C2.abs()).max()) * 2 / u
    if bound < 2 ** 24:
        assert same_bits(C1 + C2, C)
    else:   # outside the exact regime: the stated norm-wise tolerance
        assert float((C1.double() + C2.double() - C.double()).abs().max()) <= 1e-6 * bound * u


@pytest.mark.parametrize("M", [3, 4, 5])
def test_fullsize_qamaa_properties(fq, M):
    """quantize_after_mult_and_add at full size: Q(sum_k Q(a_k b_k)), table kernel."""
    xq, wq, bA, bB, exact_max = operands(fq, M)
    mv = torch.tensor([exact_max])
    mant = torch.tensor([float(M)])
    C = fq.ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1, x_bias=bA, y_bias=bB)
    assert fq.ops.last_path().startswith("qamaa-lut"), fq.ops.last_path()
    assert float(C.abs().max()) > 0
    rng = np.random.default_rng(10 + M)
    rows = np.sort(rng.choice(ROWS, size=12, replace=False))
    want = orc.qamaa_matmul(xq[rows].cpu().numpy(), np.ascontiguousarray(wq.t().cpu().numpy()), exact_max, 8, float(M), 1)
    assert same_bits(C[rows].cpu().numpy(), want)
    # arithmetic kernel (no operand biases -> no table) over the whole output
    Cg = fq.ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1)
    assert not fq.ops.last_path().startswith("qamaa-lut")
    assert same_bits(C, Cg)
    g = torch.Generator(device="cuda").manual_seed(2)
    pr = torch.randperm(ROWS, device="cuda", generator=g)
    assert same_bits(fq.ops.qamaa_matmul(xq[pr].contiguous(), wq.t(), 8, mv, mant, 1, x_bias=bA, y_bias=bB), C[pr])
    pc = torch.randperm(N_DIM, device="cuda", generator=g)
    assert same_bits(fq.ops.qamaa_matmul(xq, wq[pc].contiguous().t(), 8, mv, mant, 1, x_bias=bA,
                                         y_bias=bB[pc].contiguous()), C[:, pc])


def test_output_beyond_int32_indexing(fq):
    """Maximum sizes: an output of more than 2^31 elements (2^21 + 77 rows x 1030 columns, 8.6 GB) -- every
    index on the path must be 64-bit.  Rows at both ends and around 2^20 / 2^21 against the oracle, and the
    checksum of the rest against a second launch over a row-shifted view (row independence)."""
    E, M = 4, 3
    Mr, K, N = (1 << 21) + 77, 40, 1030
    bA, bR = 9, 7
    rng = np.random.default_rng(21)
    c = np.arange(2 ** (E + M), dtype=np.int32)
    va = orc.decode(c >> M, c & (2 ** M - 1), E, M, bA)
    va = torch.from_numpy(np.concatenate([va, -va, np.zeros(64, np.float32)]).astype(np.float32)).cuda()
    g = torch.Generator(device="cuda").manual_seed(5)
    A = va[torch.randint(0, va.numel(), (Mr, K), device="cuda", generator=g)]
    bB = rng.integers(8, 12, size=N).astype(np.int32)
    W = np.stack([orc.decode(cc >> M, cc & (2 ** M - 1), E, M, int(bB[j])) * rng.choice([-1.0, 1.0], size=K)
                  for j, cc in enumerate(rng.integers(0, 2 ** (E + M), size=(N, K)))]).astype(np.float32)
    p = fq.ops.make_params(approx_kw(M))
    C = fq.ops.approx_matmul(A, torch.from_numpy(W).cuda().t(), bA, torch.from_numpy(bB).cuda(), bR, p)
    assert fq.ops.last_path().startswith("lut") and C.numel() > 2 ** 31
    rows = np.array([0, 1, (1 << 20) - 1, 1 << 20, (1 << 21) - 1, 1 << 21, Mr - 1])
    T = orc.get_error_table_NN(E, M, True, 3)
    want = orc.approx_matmul(A[rows].cpu().numpy(), np.ascontiguousarray(W.T), T, bA, bB, bR, E=E, M=M, with_approx=1,
                             with_s2nn2s_opt=1, quant_btw_mult_accu=1)
    assert same_bits(C[rows].cpu().numpy(), want)
    # rows 3.. of the same matrix as a separate problem: its row i must equal row i + 3 of the first launch
    C2 = fq.ops.approx_matmul(A[3:], torch.from_numpy(W).cuda().t(), bA, torch.from_numpy(bB).cuda(), bR, p)
    assert same_bits(C2, C[3:])


def test_empty_conv_and_depthwise(fq):
    """Empty inputs on the conv forms: batch 0 returns an empty [0, C_out] matrix without launching."""
    p = fq.ops.make_params(approx_kw(3))
    one = torch.tensor([8.0]).cuda()
    for groups, cout in ((1, 6), (4, 4)):
        x = torch.zeros(0, 4, 8, 8).cuda()
        w = torch.zeros(cout, 4 // groups, 3, 3).cuda()
        bw = torch.full((cout,), 8.0).cuda()
        pp = fq.ops.make_params(approx_kw(3), single_column_quirk=(groups == 4))
        out = fq.ops.approx_conv2d(x, w, one, bw, one, pp, (1, 1), (1, 1), (1, 1), groups)
        assert out.shape == (0, cout)
    assert fq.ops.exact_grouped_conv2d(torch.zeros(0, 4, 8, 8).cuda(), torch.zeros(4, 1, 3, 3).cuda(), (1, 1), (1, 1),
                                       (1, 1), 4).shape == (0, 4)
