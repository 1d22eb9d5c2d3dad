"""GPU parity at model level (run with `-m gpu`): ResNet-18 and MobileNetV2 built by our mirror of the
reference's graph surgery (fp8_quantization_b200/model_wrap.py) over the same seeded FP32 models,
against tests/golden/models.npz -- produced by the UNMODIFIED reference wrappers on CPU
(tests/golden/make_models.py).

Two checks per model:
  * teacher forcing: every tapped hot-path layer is fed the reference's own calibration and eval
    inputs and must reproduce the reference's outputs (tolerance 1e-6 of the output scale: the approx
    core is bit exact, BN / activation after it are plain fp32 torch on another device);
  * end to end: logits of the whole network.  A 1-ulp difference in a BN output can move one
    activation across an FP8 rounding boundary (one grid step, ~2^-M relative) in a later layer; in
    practice the logits come out bit identical, and the test asserts 1e-6 of the logit scale.
"""
import os

import numpy as np
import pytest
import torch

import model_cases as mc

pytestmark = pytest.mark.gpu
# Round 1 on B200: 0.0 for all cases (the logits are bit identical to the reference's CPU run).
END_TO_END_TOL = 1e-6
GOLD = os.path.join(os.path.dirname(__file__), "golden", "models.npz")


def qparams(E, M):
    return dict(method="fp_quantizer", act_method="fp_quantizer", n_bits=8, n_bits_act=8, per_channel_weights=True,
                weight_range_method="current_minmax", act_range_method="allminmax", quantize_input=True,
                fp8_kwargs=dict(mantissa_bits=M, set_maxval=True), custom_approx_params=mc.approx_params(E, M),
                run_method=dict(mc.RUN_APPROX))


def build(kind, size, E, M):
    from fp8_quantization_b200 import model_wrap as mw
    if kind == "resnet18":
        model = mw.QuantizedResNet(mc.fp_resnet18(), **qparams(E, M))
    else:
        model = mw.QuantizedMobileNetV2(mc.fp_mobilenetv2(mw.MobileNetV2, size), **qparams(E, M))
    return model.cuda().eval()


@pytest.fixture(params=[False, True], ids=["reference-passes", "fused"])
def fused(request):
    """The forward with the reference's pass structure (default) and with every opt-in fusion on
    (quantization.enable_fusions: input quantizer in the operand decode, cached fixed weights, +bias -> BN ->
    activation -> NCHW store in the GEMM / conv epilogue): the goldens must hold for both."""
    from fp8_quantization_b200 import quantization as Q
    prev = Q.enable_fusions(request.param)
    yield request.param
    Q.enable_fusions(True, *prev)


@pytest.mark.parametrize("case", mc.MODEL_CASES, ids=[c[0] for c in mc.MODEL_CASES])
def test_model_against_reference(case, fused):
    name, kind, size, batch, (E, M) = case
    z = np.load(GOLD)
    if f"{name}_yeval" not in z.files:
        pytest.skip(f"{name}: golden logits not generated yet (tests/golden/make_models.py {name})")
    model = build(kind, size, E, M)
    dev = lambda a: torch.from_numpy(a).cuda()
    mods = dict(model.named_modules())
    bad = []
    with torch.inference_mode():
        model.quantized()
        for lname in mc.TAPPED_LAYERS[name]:
            layer = mods[lname]
            layer.estimate_ranges()
            got_cal = layer(dev(z[f"{name}_{lname}_xcal"]))
            layer.fix_ranges()
            got_eval = layer(dev(z[f"{name}_{lname}_xeval"]))
            for tag, got in (("cal", got_cal), ("eval", got_eval)):
                want = z[f"{name}_{lname}_y{tag}"]
                got = got.contiguous().cpu().numpy()
                tol = 1e-6 * float(np.abs(want).max())
                err = float(np.abs(got.astype(np.float64) - want).max())
                if got.shape != want.shape or err > tol:
                    bad.append((lname, tag, err, tol))
        assert not bad, bad
        # end to end (fresh calibration of the whole network); where the golden file holds probes, a fixed
        # sample of EVERY hot-path layer's output is compared bit for bit on the way (no teacher forcing)
        phase, probe_bad, hooks = ["cal"], [], []
        if name in mc.PROBED_CASES:
            for li, (lname, mod) in enumerate((n, m) for n, m in mods.items() if hasattr(m, "run_forward")):
                def probe(mod, inp, outp, lname=lname, li=li):
                    flat = outp.detach().contiguous().reshape(-1)
                    got = flat[mc.probe_indices(flat.numel(), li).to(flat.device)].cpu().numpy()
                    want = z[f"{name}_{lname}_p{phase[0]}"]
                    n_bad = int(((got.view(np.uint32) != want.view(np.uint32)) & ~((got == 0) & (want == 0))).sum())
                    if n_bad:
                        probe_bad.append((phase[0], lname, n_bad, float(np.abs(got - want).max())))
                hooks.append(mod.register_forward_hook(probe))
        model.estimate_ranges()
        y_cal = model(mc.images(batch, size, 1).cuda())
        model.fix_ranges()
        phase[0] = "eval"
        y_eval = model(mc.images(batch, size, 2).cuda())
        for h in hooks:
            h.remove()
        assert not probe_bad, probe_bad[:5]
    for tag, got in (("cal", y_cal), ("eval", y_eval)):
        want = z[f"{name}_y{tag}"]
        err = float(np.abs(got.cpu().numpy().astype(np.float64) - want).max())
        scale = float(np.abs(want).max())
        print(f"{name} {tag}: end-to-end max |dlogit| = {err:.3e} (logit scale {scale:.3e})")
        assert err <= END_TO_END_TOL * scale, (name, tag, err, scale)


def test_vit_tiny_table_vs_arithmetic():
    """ViT wrappers (models/vit_quantized_approx.py mirror) on a 2-layer toy config: the reference cannot run
    this model with approx_flag (it asserts on 3-D activations), so parity is operator
    level; here the whole forward through the table kernel must equal the forward through the arithmetic
    kernel (the definition the operator fixtures pin to the reference), calibration and eval pass."""
    from fp8_quantization_b200 import ops, vit
    cfg = vit.ViTConfig(hidden_size=64, num_hidden_layers=2, num_attention_heads=4, intermediate_size=128,
                        image_size=32, patch_size=16, num_labels=10)
    outs = []
    orig = ops.make_params
    for force in (False, True):
        torch.manual_seed(3)
        model = vit.vit_quantized_approx(vit.VisionTransformerForImageClassification(cfg), **qparams(3, 4)).cuda().eval()
        ops.make_params = (lambda *a, **k: orig(*a, **{**k, "force_generic": True})) if force else orig
        try:
            with torch.inference_mode():
                model.quantized()
                model.estimate_ranges()
                y_cal = model(mc.images(4, 32, 1).cuda())
                model.fix_ranges()
                y_eval = model(mc.images(4, 32, 2).cuda())
                path = ops.last_path()
        finally:
            ops.make_params = orig
        outs.append((y_cal, y_eval, path))
    assert outs[0][2] in ("lut", "lut-splitk", "lut-streamk") and outs[1][2] == "generic"
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert float(outs[0][1].abs().max()) > 0


def test_graphed_forward_equals_eager(fused):
    """eval_harness.GraphedForward: the calibrated ResNet-18 forward replayed as a CUDA graph returns the bits
    of the eager forward, for two different batches through the same graph and for a second input shape; and the
    fused forward returns the bits of the forward with the reference's pass structure."""
    from fp8_quantization_b200 import eval_harness as eh, quantization as Q
    res = 64
    model = build("resnet18", res, 4, 3)
    eh.calibrate(model, mc.images(4, res, 1).cuda())
    fwd = eh.GraphedForward(model)
    for seed, n in ((2, 4), (3, 4), (4, 2), (5, 4)):
        x = mc.images(n, res, seed).cuda()
        with torch.inference_mode():
            want = model(x)
            prev = Q.enable_fusions(False)
            try:
                plain = model(x)
            finally:
                Q.enable_fusions(True, *prev)
        got = fwd(x)
        assert torch.equal(got, want), (seed, n)
        assert torch.equal(plain, want), (seed, n, "fused != reference passes")
    assert len(fwd._graphs) == 2
