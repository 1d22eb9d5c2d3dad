"""Reference-pinned operator parity on the shapes of BASELINE.json configs[3] (ViT-B/16, approx_v9 E3M4) and
configs[4] (BERT-base, non-approx -> tcgen05), against tests/golden/vit_ops.npz -- outputs of the UNMODIFIED
reference's QCustomLinearTorch / QuantLayerNorm run on CPU (tests/golden/make_vit_ops.py) on a 16-column, every-4th
row sample (columns and rows are independent in the reference's loop).

Tolerances, stated per test:
  * approx GEMM (quant_btw_mult_accu: every addend a multiple of u, sums exact in fp32): BIT EXACT, also after `+ bias`;
  * non-approx matmul: |C - ref| <= 1e-6 * sum_k |a_k b_k| (fp32 summation order; north_star's tolerance);
  * the same followed by the result fake-quantizer: an order-level difference can move a value across a rounding
    boundary, i.e. by ONE grid step; at most 0.1 % of the elements may do so, none by more;
  * QuantLayerNorm: torch's layer_norm on another device, 2e-6 of the output scale; quantized weights bit exact.
"""
import os

import numpy as np
import pytest
import torch

import vit_cases as vc

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "vit_ops.npz")


def qparams(rm):
    return dict(method="fp_quantizer", act_method="fp_quantizer", n_bits=8, n_bits_act=8, per_channel_weights=True,
                weight_range_method="current_minmax", act_range_method="allminmax", quantize_input=True,
                fp8_kwargs=dict(mantissa_bits=vc.M, set_maxval=True), custom_approx_params=vc.approx_params(),
                run_method=dict(rm))


def protocol(layer, x_cal, x_eval):
    with torch.inference_mode():
        layer.quantized()
        layer.estimate_ranges()
        y_cal = layer(x_cal)
        layer.fix_ranges()
        y_eval = layer(x_eval)
    rb = layer.get_res_fp_bias()
    return dict(ycal=y_cal, yeval=y_eval, bias_a=layer.get_acts_fp_bias().reshape(-1),
                bias_w=layer.get_weights_fp_bias().reshape(-1),
                bias_r=rb.reshape(-1) if rb is not None else torch.zeros(0))


def linear(K, n_out, w, b, rm):
    from fp8_quantization_b200.approx_calculation import QCustomLinearTorch
    layer = QCustomLinearTorch(K, n_out, bias=True, **qparams(rm))
    with torch.no_grad():
        layer.weight.copy_(w)
        layer.bias.copy_(b)
    return layer.cuda().eval()


def same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | ((a == 0) & (b == 0))))


@pytest.mark.parametrize("K,N", vc.SHAPES)
@pytest.mark.parametrize("width", ["sample", "full"])
def test_vit_shaped_approx_linear_vs_reference(K, N, width):
    """width=sample: the 16-column layer the reference ran.  width=full: the real ViT-B/16 layer width (the tile /
    split-K / stream-K configurations ViT-B/16 actually uses) with the 16 sampled weight rows repeated N/16 times --
    the per-tensor result range, hence bR, is then the reference run's, and every 16-column block must reproduce it."""
    from fp8_quantization_b200 import ops
    z = np.load(GOLD)
    x, w, b = vc.operands(K, N)
    n_out = vc.COLS if width == "sample" else N
    reps = n_out // vc.COLS
    layer = linear(K, n_out, w[:vc.COLS].repeat(reps, 1), b[:vc.COLS].repeat(reps), vc.RUN_APPROX)
    got = protocol(layer, x.cuda(), (x * 0.9).cuda())
    assert ops.last_path().startswith("lut"), ops.last_path()
    pre = f"vit_{K}_{N}_"
    assert np.array_equal(got["bias_a"].cpu().numpy(), z[pre + "bias_a"])
    assert np.array_equal(got["bias_r"].cpu().numpy(), z[pre + "bias_r"])
    assert np.array_equal(got["bias_w"].cpu().numpy()[:vc.COLS], z[pre + "bias_w"])
    for tag in ("ycal", "yeval"):
        y = got[tag].cpu().numpy()[::vc.ROW_STEP]
        for r in range(reps):
            assert same_bits(y[:, r * vc.COLS:(r + 1) * vc.COLS], z[pre + tag]), (tag, r)


@pytest.mark.parametrize("K,N", vc.BERT_SHAPES)
def test_bert_shaped_plain_linear_vs_reference(K, N):
    from fp8_quantization_b200 import ops
    z = np.load(GOLD)
    x, w, b = vc.operands(K, N, vc.BERT_ROWS)
    reps = N // vc.COLS
    wf, bf = w[:vc.COLS].repeat(reps, 1), b[:vc.COLS].repeat(reps)
    # (1) x @ w.T + b, not re-quantized: norm-wise 1e-6
    layer = linear(K, N, wf, bf, vc.RUN_PLAIN_NOQ)
    got = protocol(layer, x.cuda(), (x * 0.9).cuda())
    assert ops.last_path().startswith("tcgen05"), ops.last_path()
    pre = f"bertnq_{K}_{N}_"
    assert np.array_equal(got["bias_a"].cpu().numpy(), z[pre + "bias_a"])
    assert np.array_equal(got["bias_w"].cpu().numpy()[:vc.COLS], z[pre + "bias_w"])
    with torch.inference_mode():
        wq, _ = layer.get_params()
        for tag, xin in (("ycal", x), ("yeval", x * 0.9)):
            xq = layer.activation_quantizer(xin.cuda())
            bound = 1e-6 * (xq.abs().double() @ wq[:vc.COLS].abs().double().t())[::vc.ROW_STEP].cpu().numpy()
            y = got[tag].double().cpu().numpy()[::vc.ROW_STEP]
            for r in range(0, reps, max(1, reps // 4)):
                err = np.abs(y[:, r * vc.COLS:(r + 1) * vc.COLS] - z[pre + tag])
                assert np.all(err <= bound + 1e-30), (tag, r, float((err - bound).max()))
    # (2) shipped configuration: res_quantizer(x @ w.T + b); at most 0.1 % of the elements one grid step away
    layer = linear(K, N, wf, bf, vc.RUN_PLAIN)
    got = protocol(layer, x.cuda(), (x * 0.9).cuda())
    pre = f"bert_{K}_{N}_"
    assert np.array_equal(got["bias_r"].cpu().numpy(), z[pre + "bias_r"])
    for tag in ("ycal", "yeval"):
        y = got[tag].cpu().numpy()[::vc.ROW_STEP, :vc.COLS]
        want = z[pre + tag]
        diff = y != want
        assert diff.mean() <= 1e-3, (tag, float(diff.mean()))
        step = np.maximum(np.abs(want), 2.0 ** -20) * 2.0 ** (1 - vc.M)     # > one grid step at that magnitude
        assert np.all(np.abs(y - want)[diff] <= step[diff]), tag


@pytest.mark.parametrize("tag,rm", [("ln_approx", vc.RUN_APPROX), ("ln_plain", vc.RUN_PLAIN)])
def test_quant_layernorm_vs_reference(tag, rm):
    """QuantLayerNorm (approx/replace_operations_with_approx_ops.py:160-168): input and weight fake-quantizers are
    ours (bit exact), the normalisation itself is torch's F.layer_norm on the GPU."""
    from fp8_quantization_b200.model_wrap import QuantLayerNorm
    z = np.load(GOLD)
    x, w, b = vc.layernorm_operands()
    ln = QuantLayerNorm(normalized_shape=(x.shape[1],), eps=1e-12, **qparams(rm))
    with torch.no_grad():
        ln.weight.copy_(w)
        ln.bias.copy_(b)
    ln.cuda().eval()
    got = protocol(ln, x.cuda(), (x * 0.9).cuda())
    with torch.inference_mode():
        assert same_bits(ln.weight_quantizer(ln.weight).cpu().numpy(), z[f"{tag}_wq"])
    for k in ("bias_a", "bias_w", "bias_r"):
        assert np.array_equal(got[k].cpu().numpy(), z[f"{tag}_{k}"]), k
    for k in ("ycal", "yeval"):
        y, want = got[k].cpu().numpy()[::vc.LN_ROW_STEP], z[f"{tag}_{k}"]
        scale = float(np.abs(want).max())
        if rm["original_quantize_res"]:      # output passed through the result fake-quantizer: grid-step criterion
            diff = y != want
            assert diff.mean() <= 1e-3, (k, float(diff.mean()))
            assert np.all(np.abs(y - want)[diff] <= (np.maximum(np.abs(want), 2.0 ** -20) * 2.0 ** (1 - vc.M))[diff])
        else:
            assert float(np.abs(y.astype(np.float64) - want).max()) <= 2e-6 * scale, k


def test_vit_b16_forward_table_equals_arithmetic():
    """The full ViT-B/16 forward (12 layers, hidden 768, 197 tokens, batch 4, E3M4) through the table kernels must
    equal, bit for bit, the forward through the arithmetic kernel -- the definition the operator fixtures pin to the
    reference -- in the calibration pass and the eval pass (the reference cannot run this model with approx_flag:
    it asserts on 3-D activations)."""
    from fp8_quantization_b200 import ops, vit
    import model_cases as mc
    outs = []
    orig = ops.make_params
    for force in (False, True):
        torch.manual_seed(3)
        model = vit.vit_quantized_approx(vit.VisionTransformerForImageClassification(vit.ViTConfig()),
                                         **qparams(vc.RUN_APPROX)).cuda().eval()
        ops.make_params = (lambda *a, **k: orig(*a, **{**k, "force_generic": True})) if force else orig
        try:
            with torch.inference_mode():
                model.quantized()
                model.estimate_ranges()
                y_cal = model(mc.images(4, 224, 1).cuda())
                model.fix_ranges()
                y_eval = model(mc.images(4, 224, 2).cuda())
        finally:
            ops.make_params = orig
        outs.append((y_cal, y_eval))
        del model
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert float(outs[0][1].abs().max()) > 0
