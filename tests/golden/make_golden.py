"""Generates the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE (CPU).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The fixtures are what travels to the GPU box; the reference does not.

Fixtures written (all npz, small):
  pairs_E{E}M{M}.npz      all signed codes x all signed codes through custom_matmul_vectorize
                          (approx/approx_matmul_whole_v9.py:10-169) for a set of flag/bias combos,
                          python-int biases and int32-tensor biases (single-column quirk branch)
  debug_params.npz        the reference's own dumped MobileNetV2 operands (debug_params/*.csv) and
                          the approx_multiply column loop output (approx_calculation.py:946-971)
  gemm_random.npz         value-space-sampled and off-grid random GEMMs
  fakequant.npz           quantize_to_fp8_ste_MM (quantization/quantizers/fp8_quantizer.py:97-173)
  codec.npz               float_to_fpany_absint_torch / fpany_absint_to_float_torch / allnorm codec
"""
import itertools
import os
import sys
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402

warnings.filterwarnings("ignore")
torch.set_num_threads(os.cpu_count())

FORMATS = [(4, 3), (3, 4), (2, 5)]
# (withComp, dnsmp_factor) choices that select distinct tables per format (v9 :555-592)
TABLE_CHOICES = {
    (4, 3): [(True, 3), (False, 3)],
    (3, 4): [(True, 3), (True, 4), (False, 3)],
    (2, 5): [(True, 3), (True, 4), (True, 5), (False, 3)],
}


def all_codes(v9, E, M, bias, extra_top=True):
    """All non-negative grid values of the format (+ one binade above max_expo, as the FP8
    fake-quantizer can emit when maxval rounds up), then both signs."""
    pd = v9.param_prepare(E, M, custom_bias=bias)
    n = 2 ** (E + M) + (2 ** M if extra_top else 0)
    codes = torch.arange(n, dtype=torch.int32)
    vals = v9.fpany_absint_to_float_torch(pd, abs_int=codes).float()
    return torch.cat([vals, -vals])


def run_v9(v9, A, B, E, M, bA, bB, bR, withComp, dnsmp, flags, tensor_bias=False):
    table = v9.get_error_table_NN(E, M, withComp, dnsmp)
    if tensor_bias:
        bA = torch.tensor([bA], dtype=torch.int32)
        bB = torch.tensor(bB, dtype=torch.int32)
        bR = torch.tensor([bR], dtype=torch.int32)
    return v9.custom_matmul_vectorize(A, B, E, M, bA, bB, bR, table, **flags)


def flag_dict(with_approx, s2nn2s, qbma, clip):
    return dict(with_approx=bool(with_approx), with_s2nn2s_opt=bool(s2nn2s),
                golden_clip_OF=bool(clip), quant_btw_mult_accu=bool(qbma))


def gen_pairs(v9):
    for (E, M) in FORMATS:
        half = 2 ** (E - 1)
        out = {}
        cfgs = []
        bias_sets = [(half, half, half), (half + 1, half + 3, half + 2), (half + 2, half + 1, half - 1)]
        picks = [  # (wa, s2, qb, clip)
            (1, 1, 1, 0), (1, 0, 1, 0), (1, 1, 0, 0), (1, 0, 0, 0), (0, 1, 1, 1), (1, 1, 1, 1), (0, 0, 1, 0),
        ]
        idx = 0
        for (wc, dn) in TABLE_CHOICES[(E, M)]:
            for bi, (bA, bB, bR) in enumerate(bias_sets):
                for fl in picks:
                    for tb in (False, True):
                        # keep the fixture small: full flag set only on the first bias set
                        if bi > 0 and fl not in picks[:2]:
                            continue
                        A = all_codes(v9, E, M, bA).reshape(-1, 1)
                        B = all_codes(v9, E, M, bB).reshape(1, -1)
                        res = run_v9(v9, A, B, E, M, bA, bB, bR, wc, dn, flag_dict(*fl), tensor_bias=tb)
                        out[f"out{idx}"] = res.numpy().astype(np.float32)
                        cfgs.append([E, M, int(wc), dn, bA, bB, bR, *fl, int(tb)])
                        idx += 1
        out["cfgs"] = np.array(cfgs, dtype=np.int32)
        out["cfg_cols"] = np.array(["E", "M", "withComp", "dnsmp", "bA", "bB", "bR", "with_approx",
                                    "s2nn2s", "qbma", "clip_OF", "tensor_bias"])
        np.savez_compressed(os.path.join(HERE, f"pairs_E{E}M{M}.npz"), **out)
        print(f"pairs E{E}M{M}: {idx} configs")


def gen_debug_params(v9):
    d = os.path.join(ref_shim.REF_ROOT, "debug_params")
    act = np.loadtxt(os.path.join(d, "act.csv"), delimiter=",", dtype=np.float32)
    wgt = np.loadtxt(os.path.join(d, "weight.csv"), delimiter=",", dtype=np.float32)
    wb = np.loadtxt(os.path.join(d, "weight_bias.csv"), delimiter=",", dtype=np.float32).astype(np.int32)
    ab = int(np.loadtxt(os.path.join(d, "act_bias.csv"), delimiter=",", dtype=np.float32))
    E, M = 3, 4
    A = torch.from_numpy(act)
    B = torch.from_numpy(wgt)
    out = {"act": act, "weight": wgt, "weight_bias": wb.reshape(-1), "act_bias": np.int32(ab)}
    cfgs = []
    idx = 0
    for (wc, dn) in TABLE_CHOICES[(E, M)]:
        for bR in (4, 6, 8):
            for fl in [(1, 1, 1, 0), (1, 0, 1, 0), (1, 1, 0, 0), (1, 0, 0, 0)]:
                if bR != 6 and fl != (1, 1, 1, 0):
                    continue
                cols = []
                for i in range(B.shape[1]):   # approx_calculation.py:948-958
                    cols.append(run_v9(v9, A, B[:, i:i + 1], E, M, ab, int(wb.reshape(-1)[i]), bR, wc, dn,
                                       flag_dict(*fl)))
                out[f"out{idx}"] = torch.cat(cols, dim=1).numpy()
                cfgs.append([E, M, int(wc), dn, ab, -1, bR, *fl, 0])
                idx += 1
    out["cfgs"] = np.array(cfgs, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "debug_params.npz"), **out)
    print("debug_params:", idx, "configs", act.shape, wgt.shape)


def sample_grid(v9, E, M, bias, shape, rng, zero_frac=0.2):
    vals = all_codes(v9, E, M, bias, extra_top=False).numpy()
    x = rng.choice(vals, size=shape).astype(np.float32)
    x[rng.random(shape) < zero_frac] = 0.0
    return x


def gen_gemm_random(v9):
    rng = np.random.default_rng(5)
    out = {}
    cfgs = []
    idx = 0
    for (E, M) in FORMATS:
        half = 2 ** (E - 1)
        for (Mr, K, N) in [(33, 9, 5), (17, 144, 8), (8, 576, 4)]:
            bA, bR = half + 1, half
            bB = rng.integers(half, half + 3, size=N).astype(np.int32)
            A = sample_grid(v9, E, M, bA, (Mr, K), rng)
            B = np.stack([sample_grid(v9, E, M, int(bB[j]), (K,), rng, 0.05) for j in range(N)], axis=1)
            # scale weights down so sums stay in range
            for (wc, dn) in TABLE_CHOICES[(E, M)][:2]:
                for fl in [(1, 1, 1, 0), (1, 0, 0, 0)]:
                    cols = [run_v9(v9, torch.from_numpy(A), torch.from_numpy(B[:, j:j + 1].copy()), E, M, bA,
                                   int(bB[j]), bR, wc, dn, flag_dict(*fl)) for j in range(N)]
                    out[f"A{idx}"], out[f"B{idx}"], out[f"bB{idx}"] = A, B, bB
                    out[f"out{idx}"] = torch.cat(cols, dim=1).numpy()
                    cfgs.append([E, M, int(wc), dn, bA, -1, bR, *fl, 0])
                    idx += 1
        # off-grid operands (plain gaussians): exercises the rounding paths of the encoder
        A = (rng.standard_normal((16, 24)) * 0.5).astype(np.float32)
        B = (rng.standard_normal((24, 6)) * 0.1).astype(np.float32)
        bB = np.full(6, half + 2, np.int32)
        for fl in [(1, 1, 1, 0), (1, 0, 1, 1), (1, 1, 0, 0), (0, 0, 0, 0)]:
            wc, dn = TABLE_CHOICES[(E, M)][0]
            cols = [run_v9(v9, torch.from_numpy(A), torch.from_numpy(B[:, j:j + 1].copy()), E, M, half, int(bB[j]),
                           half, wc, dn, flag_dict(*fl)) for j in range(6)]
            out[f"A{idx}"], out[f"B{idx}"], out[f"bB{idx}"] = A, B, bB
            out[f"out{idx}"] = torch.cat(cols, dim=1).numpy()
            cfgs.append([E, M, int(wc), dn, half, -1, half, *fl, 0])
            idx += 1
    out["cfgs"] = np.array(cfgs, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "gemm_random.npz"), **out)
    print("gemm_random:", idx, "configs")


def gen_fakequant(ns):
    q = ns.fp8q.quantize_to_fp8_ste_MM
    rng = np.random.default_rng(11)
    out = {}
    cfgs = []
    idx = 0
    for mbits in (2, 3, 4, 5):
        # per tensor, gaussian + edge values (0, +-maxval, neighbours of powers of two)
        x = (rng.standard_normal(4096) * 1.3).astype(np.float32)
        p2 = np.float32(2.0) ** np.arange(-12, 4, dtype=np.float32)
        edges = np.concatenate([p2, np.nextafter(p2, np.float32(0)), np.nextafter(p2, np.float32(9)),
                                -p2, [0.0, -0.0]]).astype(np.float32)
        x = np.concatenate([x, edges])
        for maxval in (np.float32(np.abs(x).max()), np.float32(3.0), np.float32(0.7311), np.float32(6.0)):
            xx = np.concatenate([x, [maxval, -maxval]]).astype(np.float32)
            y, b = q(torch.from_numpy(xx), 8, torch.tensor([maxval]), torch.tensor([float(mbits)]), 1)
            out[f"x{idx}"], out[f"mv{idx}"], out[f"y{idx}"], out[f"b{idx}"] = xx, np.array([maxval]), y.numpy(), b.numpy()
            cfgs.append([mbits, 1, 0])
            idx += 1
        # unsigned (sign_bits = 0) branch, fp8_quantizer.py:119
        xu = np.abs(x)
        y, b = q(torch.from_numpy(xu), 8, torch.tensor([np.float32(2.5)]), torch.tensor([float(mbits)]), 0)
        out[f"x{idx}"], out[f"mv{idx}"], out[f"y{idx}"], out[f"b{idx}"] = xu, np.array([2.5], np.float32), y.numpy(), b.numpy()
        cfgs.append([mbits, 0, 0])
        idx += 1
        # per channel weights [C,Cin,kh,kw] and [N,K]
        for shape in [(8, 4, 3, 3), (16, 40)]:
            w = (rng.standard_normal(shape) * rng.uniform(0.01, 0.5, size=(shape[0],) + (1,) * (len(shape) - 1))
                 ).astype(np.float32)
            mv = np.abs(w.reshape(shape[0], -1)).max(axis=1).astype(np.float32)
            y, b = q(torch.from_numpy(w), 8, torch.from_numpy(mv), torch.tensor([float(mbits)]), 1)
            out[f"x{idx}"], out[f"mv{idx}"], out[f"y{idx}"], out[f"b{idx}"] = w, mv, y.numpy(), b.numpy()
            cfgs.append([mbits, 1, 1])
            idx += 1
    out["cfgs"] = np.array(cfgs, dtype=np.int32)  # mantissa_bits, sign_bits, per_channel
    np.savez_compressed(os.path.join(HERE, "fakequant.npz"), **out)
    print("fakequant:", idx, "configs")


def gen_codec(ns):
    v9, v5 = ns.v9, ns.fp_any_v5
    rng = np.random.default_rng(3)
    out = {}
    cfgs = []
    idx = 0
    for (E, M) in FORMATS + [(5, 2)]:
        for bias in (2 ** (E - 1) - 1, 2 ** (E - 1) + 2):
            pd = v9.param_prepare(E, M, custom_bias=bias)
            grid = all_codes(v9, E, M, bias).numpy()
            x = np.concatenate([grid, (rng.standard_normal(2048) * pd["max_norm"] / 3).astype(np.float32),
                                (rng.standard_normal(1024) * pd["min_norm"]).astype(np.float32)]).astype(np.float32)
            for clip in (False, True):
                e, m = v9.float_to_fpany_absint_torch(pd, torch.from_numpy(x), clip_OF=clip)
                qv = v9.quant_to_fp_any_vectorize_torch(torch.from_numpy(x), E, M, custom_bias=bias, clip_OF=clip)
                pd5 = v5.param_prepare(E, M, custom_bias=bias)
                e5, m5 = v5.float_to_fpany_absint_torch_allnorm(pd5, torch.from_numpy(x), clip_OF=clip)
                d5 = v5.fpany_absint_to_float_torch_allnorm(pd5, expo=e5, mant=m5)
                out[f"x{idx}"] = x
                out[f"e{idx}"], out[f"m{idx}"], out[f"q{idx}"] = e.numpy(), m.numpy(), qv.numpy()
                out[f"e5_{idx}"], out[f"m5_{idx}"], out[f"d5_{idx}"] = e5.numpy(), m5.numpy(), d5.numpy()
                cfgs.append([E, M, bias, int(clip)])
                idx += 1
    out["cfgs"] = np.array(cfgs, dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "codec.npz"), **out)
    print("codec:", idx, "configs")


def main():
    ns = ref_shim.load()
    which = sys.argv[1:] or ["pairs", "debug", "gemm", "fakequant", "codec"]
    with torch.inference_mode():
        if "pairs" in which:
            gen_pairs(ns.v9)
        if "debug" in which:
            gen_debug_params(ns.v9)
        if "gemm" in which:
            gen_gemm_random(ns.v9)
        if "fakequant" in which:
            gen_fakequant(ns)
        if "codec" in which:
            gen_codec(ns)


if __name__ == "__main__":
    main()
