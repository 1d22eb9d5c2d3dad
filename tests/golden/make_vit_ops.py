"""Generates tests/golden/vit_ops.npz by RUNNING THE UNMODIFIED REFERENCE on CPU (build container only):

  QCustomLinearTorch   approx/approx_calculation.py:920-1023   ViT-B/16-shaped GEMMs (tests/vit_cases.py), approx_v9
                                                               E3M4, a 16-column sample of each layer
  QuantLayerNorm       approx/replace_operations_with_approx_ops.py:160-168   [3152, 768]
  QCustomLinearTorch   non-approx configuration (`x @ y[:, i]` per column, :968-969) on BERT-base shapes
                       4096 x 768 x {768, 3072}, 3072 -> 768: BASELINE configs[4], 16-column sample

each through quantized() -> estimate_ranges() -> forward -> fix_ranges() -> forward.  Inputs are regenerated from
seeds by the tests (tests/vit_cases.py); stored: outputs and the fp biases.

    python tests/golden/make_vit_ops.py
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
import ref_shim  # noqa: E402
import vit_cases as vc  # noqa: E402

warnings.filterwarnings("ignore")
torch.set_num_threads(os.cpu_count())


def qparams(run_method):
    from quantization.quantizers import FPQuantizer
    from quantization.range_estimators import RangeEstimators
    return dict(method=FPQuantizer, act_method=FPQuantizer, n_bits=8, n_bits_act=8, per_channel_weights=True,
                weight_range_method=RangeEstimators.current_minmax.cls, act_range_method=RangeEstimators.allminmax.cls,
                quantize_input=True, fp8_kwargs=dict(mantissa_bits=vc.M, set_maxval=True),
                custom_approx_params=vc.approx_params(), run_method=dict(run_method))


def protocol(layer, x_cal, x_eval, step):
    with torch.inference_mode(), contextlib.redirect_stdout(io.StringIO()):
        layer.quantized()
        layer.estimate_ranges()
        y_cal = layer(x_cal)
        layer.fix_ranges()
        y_eval = layer(x_eval)
    rb = layer.get_res_fp_bias()
    # rows are independent: a fixed row sample (vit_cases.ROW_STEP) keeps the fixture small
    return dict(ycal=y_cal.contiguous().numpy()[::step], yeval=y_eval.contiguous().numpy()[::step],
                bias_a=layer.get_acts_fp_bias().reshape(-1).numpy(), bias_w=layer.get_weights_fp_bias().reshape(-1).numpy(),
                bias_r=rb.reshape(-1).numpy() if rb is not None else np.zeros(0, np.float32))


def main():
    ac = ref_shim.load_operators()
    import importlib
    rep = importlib.import_module("approx.replace_operations_with_approx_ops")
    out = {}
    for tag, shapes, rows, rm in (("vit", vc.SHAPES, vc.ROWS, vc.RUN_APPROX), ("bert", vc.BERT_SHAPES, vc.BERT_ROWS, vc.RUN_PLAIN),
                                 ("bertnq", vc.BERT_SHAPES, vc.BERT_ROWS, vc.RUN_PLAIN_NOQ)):
        for K, N in shapes:
            t0 = time.time()
            x, w, b = vc.operands(K, N, rows)
            layer = ac.QCustomLinearTorch(K, vc.COLS, bias=True, **qparams(rm))
            with torch.no_grad():
                layer.weight.copy_(w[:vc.COLS])
                layer.bias.copy_(b[:vc.COLS])
            layer.eval()
            for k, v in protocol(layer, x, x * 0.9, vc.ROW_STEP).items():
                out[f"{tag}_{K}_{N}_{k}"] = v
            print(tag, K, N, "done in", round(time.time() - t0, 1), "s", flush=True)
    x, w, b = vc.layernorm_operands()
    for tag, rm in (("ln_approx", vc.RUN_APPROX), ("ln_plain", vc.RUN_PLAIN)):
        ln = rep.QuantLayerNorm(normalized_shape=(x.shape[1],), eps=1e-12, **qparams(rm))
        with torch.no_grad():
            ln.weight.copy_(w)
            ln.bias.copy_(b)
        ln.eval()
        for k, v in protocol(ln, x, x * 0.9, vc.LN_ROW_STEP).items():
            out[f"{tag}_{k}"] = v
        with torch.inference_mode():
            out[f"{tag}_wq"] = ln.weight_quantizer(ln.weight).numpy()
    np.savez_compressed(os.path.join(HERE, "vit_ops.npz"), **out)
    print("vit_ops:", len(out), "arrays")


if __name__ == "__main__":
    main()
