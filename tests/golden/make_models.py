"""Generates tests/golden/models.npz by running the UNMODIFIED reference model wrappers on CPU:
  models/resnet_quantized_approx.py:46-130      QuantizedResNet (torchvision resnet18, random init)
  models/mobilenet_v2_quantized_approx.py:26-88 QuantizedMobileNetV2 (models/mobilenet_v2.py)
with the scripts' quantizer flags (--fp8-set-maxval --per-channel --quantize-input, act allminmax,
weight current_minmax) and approx_v9 (withComp, with_approx, with_s2nn2s_opt, quant_btw_mult_accu),
through estimate_ranges -> one calibration batch -> fix_ranges -> eval batch.  Stored: both input
batches and both logits tensors (weights are regenerated from seeds by tests/model_cases.py).

Build container only:   python tests/golden/make_models.py [case ...]
"""
import contextlib
import io
import os
import sys
import time
import warnings

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
import model_cases as mc  # noqa: E402
import ref_shim  # noqa: E402

warnings.filterwarnings("ignore")
torch.set_num_threads(os.cpu_count())


def qparams(E, M):
    from quantization.quantizers import FPQuantizer
    from quantization.range_estimators import RangeEstimators
    return dict(method=FPQuantizer, act_method=FPQuantizer, n_bits=8, n_bits_act=8, per_channel_weights=True,
                weight_range_method=RangeEstimators.current_minmax.cls, act_range_method=RangeEstimators.allminmax.cls,
                quantize_input=True, fp8_kwargs=dict(mantissa_bits=M, set_maxval=True),
                custom_approx_params=mc.approx_params(E, M), run_method=dict(mc.RUN_APPROX))


def main():
    ref_shim.load_operators()
    # models/__init__.py imports every architecture (ViT needs an older transformers, BERT downloads
    # weights at import time): register an empty package that only provides the search path.
    import types
    pkg = types.ModuleType("models")
    pkg.__path__ = [os.path.join(ref_shim.REF_ROOT, "models")]
    sys.modules["models"] = pkg
    from models.resnet_quantized_approx import QuantizedResNet
    from models.mobilenet_v2_quantized_approx import QuantizedMobileNetV2
    from models.mobilenet_v2 import MobileNetV2
    path = os.path.join(HERE, "models.npz")
    out = dict(np.load(path)) if os.path.exists(path) else {}
    want = sys.argv[1:]
    for name, kind, size, batch, (E, M) in mc.MODEL_CASES:
        if want and name not in want:
            continue
        if kind == "resnet18":
            model = QuantizedResNet(mc.fp_resnet18(), **qparams(E, M))
        else:
            model = QuantizedMobileNetV2(mc.fp_mobilenetv2(MobileNetV2, size), **qparams(E, M))
        model.eval()
        x_cal, x_eval = mc.images(batch, size, 1), mc.images(batch, size, 2)
        t0 = time.time()
        # per-layer taps (teacher forcing on our side): input and output of selected hot-path layers
        taps, phase = {}, ["cal"]
        mods = dict(model.named_modules())
        for lname in mc.TAPPED_LAYERS[name]:
            def hook(mod, inp, outp, lname=lname):
                taps[f"{name}_{lname}_x{phase[0]}"] = inp[0].detach().contiguous().numpy().copy()
                taps[f"{name}_{lname}_y{phase[0]}"] = outp.detach().contiguous().numpy().copy()
            mods[lname].register_forward_hook(hook)
        # probes: for EVERY hot-path layer a fixed pseudo-random sample of its output (end-to-end pass, no
        # teacher forcing) -- locates the first layer at which two implementations part ways
        if name in mc.PROBED_CASES:
            for li, (lname, mod) in enumerate((n, m) for n, m in mods.items() if hasattr(m, "run_forward")):
                def probe(mod, inp, outp, lname=lname, li=li):
                    flat = outp.detach().contiguous().reshape(-1)
                    idx = mc.probe_indices(flat.numel(), li)
                    taps[f"{name}_{lname}_p{phase[0]}"] = flat[idx].numpy().copy()
                mod.register_forward_hook(probe)
        with torch.inference_mode(), contextlib.redirect_stdout(io.StringIO()):
            model.quantized()
            model.estimate_ranges()
            y_cal = model(x_cal)
            model.fix_ranges()
            phase[0] = "eval"
            y_eval = model(x_eval)
        out.update(taps)
        if x_cal.numel() < 1 << 16:      # large inputs are regenerated from their seeds by the test
            out[f"{name}_xcal"], out[f"{name}_xeval"] = x_cal.numpy(), x_eval.numpy()
        out[f"{name}_ycal"], out[f"{name}_yeval"] = y_cal.numpy(), y_eval.numpy()
        print(name, "done in", round(time.time() - t0, 1), "s; logits abs max", float(y_eval.abs().max()), flush=True)
        np.savez_compressed(path, **out)


if __name__ == "__main__":
    main()
