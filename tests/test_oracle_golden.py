"""Pins the CPU oracle (oracle/approx_oracle.c) to the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import oracle as orc

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def signed_codes(E, M, bias, extra_top=True):
    n = 2 ** (E + M) + (2 ** M if extra_top else 0)
    c = np.arange(n, dtype=np.int32)
    v = orc.decode(c >> M, c & (2 ** M - 1), E, M, bias)
    return np.concatenate([v, -v]).astype(np.float32)


def cfg_kwargs(row):
    E, M, wc, dn, bA, bB, bR, wa, s2, qb, clip, tb = [int(v) for v in row]
    kw = dict(E=E, M=M, with_approx=wa, with_s2nn2s_opt=s2, quant_btw_mult_accu=qb, golden_clip_OF=clip,
              tensor_bias_quirk=tb)
    return kw, orc.get_error_table_NN(E, M, bool(wc), dn), bA, bB, bR


def same_bits(a, b):
    a = np.asarray(a, np.float32)
    b = np.asarray(b, np.float32)
    both_nan = np.isnan(a) & np.isnan(b)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | both_nan | ((a == 0) & (b == 0))))


@pytest.mark.parametrize("fmt", [(4, 3), (3, 4), (2, 5)])
def test_exhaustive_pairs(fmt):
    """Every signed code x every signed code, K=1 (no summation-order effect): bit exact."""
    E, M = fmt
    z = np.load(os.path.join(GOLD, f"pairs_E{E}M{M}.npz"))
    cfgs = z["cfgs"]
    bad = []
    for i, row in enumerate(cfgs):
        kw, T, bA, bB, bR = cfg_kwargs(row)
        a = signed_codes(E, M, bA).reshape(-1, 1)
        b = signed_codes(E, M, bB).reshape(1, -1)
        got = orc.pair(a, b, T, bA, bB, bR, **kw)
        if not same_bits(got, z[f"out{i}"]):
            d = np.argwhere(got.view(np.uint32) != z[f"out{i}"].view(np.uint32))
            bad.append((i, row.tolist(), len(d), d[:3].tolist()))
    assert not bad, bad[:5]


def _check_gemm(got, ref, absum, unit_exact_ok=True):
    """Exact where fp32 summation is order independent, else norm-wise 1e-6 (stated tolerance)."""
    tol = 1e-6 * absum + 1e-30
    assert np.all(np.abs(got.astype(np.float64) - ref.astype(np.float64)) <= tol)


def test_debug_params_fixture():
    z = np.load(os.path.join(GOLD, "debug_params.npz"))
    A, B, bB, bA = z["act"], z["weight"], z["weight_bias"], int(z["act_bias"])
    exact = 0
    for i, row in enumerate(z["cfgs"]):
        kw, T, _, _, bR = cfg_kwargs(row)
        got = orc.approx_matmul(A, B, T, bA, bB, bR, **kw)
        ab = orc.approx_matmul_abs(A, B, T, bA, bB, bR, **kw)
        _check_gemm(got, z[f"out{i}"], ab)
        if kw["quant_btw_mult_accu"]:
            # addends are multiples of 2^(1-bR-M) and sums stay below 2^24 units: order free
            assert same_bits(got, z[f"out{i}"]), row
            exact += 1
    assert exact >= 10


def test_gemm_random_fixture():
    z = np.load(os.path.join(GOLD, "gemm_random.npz"))
    nexact = 0
    for i, row in enumerate(z["cfgs"]):
        kw, T, bA, _, bR = cfg_kwargs(row)
        A, B, bB = z[f"A{i}"], z[f"B{i}"], z[f"bB{i}"]
        got = orc.approx_matmul(A, B, T, bA, bB, bR, **kw)
        ab = orc.approx_matmul_abs(A, B, T, bA, bB, bR, **kw)
        _check_gemm(got, z[f"out{i}"], ab)
        nexact += same_bits(got, z[f"out{i}"])
    assert nexact >= len(z["cfgs"]) // 2


def test_fakequant_fixture():
    z = np.load(os.path.join(GOLD, "fakequant.npz"))
    for i, (mbits, sbits, pc) in enumerate(z["cfgs"]):
        y, b = orc.fp8_fakequant(z[f"x{i}"], z[f"mv{i}"], 8, float(mbits), int(sbits))
        assert np.array_equal(b.reshape(-1), z[f"b{i}"].reshape(-1)), (i, b, z[f"b{i}"])
        assert same_bits(y, z[f"y{i}"]), i


def same_bits_or_nan(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0)) | (np.isnan(a) & np.isnan(b))))


def test_fakequant_special_values():
    """NaN propagates (torch.min / torch.max), +-Inf clamps to +-maxval, sub-normals and signed zeros: reference outputs."""
    z = np.load(os.path.join(GOLD, "fakequant_special.npz"))
    for i, (mbits, sbits) in enumerate(z["cfgs"]):
        y, b = orc.fp8_fakequant(z["x"], z[f"mv{i}"], 8, float(mbits), int(sbits))
        assert np.array_equal(b.reshape(-1), z[f"b{i}"].reshape(-1)), i
        assert same_bits_or_nan(y, z[f"y{i}"]), (i, y, z[f"y{i}"])
    assert np.isnan(z["y0"]).sum() > 0


def test_codec_fixture():
    z = np.load(os.path.join(GOLD, "codec.npz"))
    for i, (E, M, bias, clip) in enumerate(z["cfgs"]):
        E, M, bias, clip = int(E), int(M), int(bias), int(clip)
        x = z[f"x{i}"]
        e, m = orc.encode(x, E, M, bias, clip_OF=clip)
        assert np.array_equal(e, z[f"e{i}"]) and np.array_equal(m, z[f"m{i}"])
        assert same_bits(orc.quant(x, E, M, bias, clip_OF=clip), z[f"q{i}"])
        e5, m5 = orc.allnorm_encode(x, E, M, bias, clip_OF=clip)
        assert np.array_equal(e5, z[f"e5_{i}"]) and np.array_equal(m5, z[f"m5_{i}"])
        assert same_bits(orc.allnorm_decode(e5, m5, M, bias), z[f"d5_{i}"])


def test_grid_fixed_points():
    """Known-answer property from the reference's test.py:59-71: every grid value is a fixed point
    of the fake-quantizer; and decode(encode(v)) == v for all codes (fp_any_int_v5.py:214-221)."""
    for (E, M) in [(4, 3), (3, 4), (2, 5), (5, 2)]:
        for bias in (2 ** (E - 1), 2 ** (E - 1) + 2):
            v = signed_codes(E, M, bias, extra_top=False)
            e, m = orc.encode(np.abs(v), E, M, bias)
            assert same_bits(orc.decode(e, m, E, M, bias), np.abs(v))
            maxval = float(np.abs(v).max())
            y, b = orc.fp8_fakequant(v, [maxval], 8, float(M), 1)
            assert b[0] == bias
            assert same_bits(y, v)


def test_v5_ofuf_fixture():
    """Row A7: the oracle's overflow / underflow clipping against the unmodified reference's
    approx_mult_new (approx/approx_matmul_whole_v5.py:155-182), all 8 flag combinations."""
    z = np.load(os.path.join(GOLD, "v5_ofuf.npz"))
    assert int(z["of_opt_raises_nameerror"]) == 1     # the reference line :173 as shipped
    for i, (E, M, wc, dn, B_neg, sim, of, uf, g) in enumerate(z["cfgs"]):
        x, y, T = z[f"x{g}"], z[f"y{g}"], z[f"t{g}"].astype(np.int32)
        comp = T[x % (2 ** M), y % (2 ** M)]
        got = orc.v5_ofuf(x, y, comp, int(B_neg), int(E), int(M), int(sim), int(of), int(uf))
        assert np.array_equal(got, z[f"r{i}"]), (i, E, M, sim, of, uf)


def test_v5_gemm_fixture():
    """Oracle restatement of the v5 GEMM (approx/approx_matmul_whole_v5.py:10-182, the live OF/UF flags) against
    outputs of the unmodified reference function (tests/golden/make_v5_gemm.py): bit exact with sim_hw_add_OFUF
    (addends are bounded multiples of u: the fp32 sum is order independent), 1e-6 norm-wise otherwise."""
    z = np.load(os.path.join(GOLD, "v5_gemm.npz"))
    nexact = 0
    for i, (E, M, bias, ongrid, wc, sim, of, uf, g) in enumerate(z["cfgs"].tolist()):
        T = orc.v5_comp_table(E, M, bool(wc), 3)
        A, B = z[f"A{g}"], z[f"B{g}"]
        got = orc.v5_matmul(A, B, T, E, M, bias, bias, bias, sim, of, uf)
        want = z[f"r{i}"]
        if same_bits(got, want):
            nexact += 1
            continue
        assert not sim, (i, "not bit exact with sim_hw_add_OFUF")
        mag = np.abs(orc.v5_matmul(np.abs(A), np.abs(B), T, E, M, bias, bias, bias, sim, of, uf, sum_mode=1)).astype(np.float64)
        assert np.all(np.abs(got.astype(np.float64) - want) <= 1e-6 * mag + 1e-30), i
    assert nexact >= len(z["cfgs"]) * 3 // 4


def test_v5_comp_tables_match_reference():
    import fp8_quantization_b200 as pkg
    z = np.load(os.path.join(GOLD, "comp_tables_v5.npz"))
    want = {(4, 3, 3): 0, (3, 4, 3): 1, (3, 4, 4): 2, (3, 4, 5): 2, (2, 5, 3): 3, (2, 5, 4): 4, (2, 5, 5): 5, (2, 5, 7): 5}
    for (E, M, d), idx in want.items():
        assert np.array_equal(pkg.ops.comp_table_v5(E, M, True, d).numpy(), z[f"t{idx}"]), (E, M, d)
        assert np.array_equal(orc.v5_comp_table(E, M, True, d), z[f"t{idx}"])
        assert not pkg.ops.comp_table_v5(E, M, False, d).any()
    with pytest.raises(UnboundLocalError):
        pkg.ops.comp_table_v5(3, 4, True, 2)
    with pytest.raises(ValueError):
        pkg.ops.comp_table_v5(5, 2, True, 3)
