"""Generates tests/golden/layers.npz by RUNNING THE UNMODIFIED REFERENCE operator classes on CPU:

  QCustomLinearTorch    approx/approx_calculation.py:920-1023
  QCustomBNConv2dTorch  approx/approx_calculation.py:723-917   (groups = 1, grouped, depthwise)

through the reference's own protocol (quantization/hijacker.py:77-115,
quantization/quantized_folded_bn.py:30-83): quantized() -> estimate_ranges() -> forward(calibration
batch) -> fix_ranges() -> forward(eval batch).  Stored per case: constructor arguments, weights,
inputs, the calibration-pass output, the eval output and the three fp biases the layer ended with.

Run in the build container only (needs /root/reference):   python tests/golden/make_layers.py
"""
import contextlib
import io
import json
import os
import sys
import warnings

import numpy as np
import torch
from torch import nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_shim  # noqa: E402

warnings.filterwarnings("ignore")
torch.set_num_threads(os.cpu_count())


def approx_params(E, M, **kw):
    d = dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True, with_s2nn2s_opt=True,
             sim_hw_add_OFUF=False, with_OF_opt=False, with_UF_opt=False, golden_clip_OF=False,
             quant_btw_mult_accu=True, debug_mode=False, self_check_mode=False)
    d.update(kw)
    return d


def run_method(approx=True, qamaa=False, res=True, orig=False):
    return dict(approx_flag=approx, quantize_after_mult_and_add=qamaa, res_quantizer_flag=res,
                original_quantize_res=orig)


# kind, ctor args, input shape, approx params, run method, activation
CASES = [
    dict(kind="linear", args=dict(in_features=96, out_features=40), x=(24, 96), ap=approx_params(4, 3), rm=run_method()),
    dict(kind="linear", args=dict(in_features=70, out_features=33), x=(17, 70), ap=approx_params(3, 4), rm=run_method()),
    dict(kind="linear", args=dict(in_features=64, out_features=24), x=(16, 64),
         ap=approx_params(2, 5, dnsmp_factor=4), rm=run_method()),
    dict(kind="linear", args=dict(in_features=64, out_features=24), x=(16, 64),
         ap=approx_params(3, 4, with_s2nn2s_opt=False, withComp=False), rm=run_method()),
    dict(kind="linear", args=dict(in_features=64, out_features=24), x=(16, 64),
         ap=approx_params(4, 3, golden_clip_OF=True), rm=run_method(orig=True)),
    dict(kind="linear", args=dict(in_features=48, out_features=20), x=(12, 48), ap=approx_params(4, 3),
         rm=run_method(approx=False, qamaa=True)),
    dict(kind="linear", args=dict(in_features=48, out_features=20), x=(12, 48), ap=approx_params(3, 4),
         rm=run_method(approx=False, qamaa=True)),
    dict(kind="linear", args=dict(in_features=48, out_features=20), x=(12, 48), ap=approx_params(3, 4),
         rm=run_method(approx=False, qamaa=False, orig=True)),
    dict(kind="conv", args=dict(in_channels=8, out_channels=12, kernel_size=3, stride=1, padding=1), x=(2, 8, 10, 10),
         ap=approx_params(4, 3), rm=run_method(), act="relu"),
    dict(kind="conv", args=dict(in_channels=6, out_channels=10, kernel_size=3, stride=2, padding=1), x=(2, 6, 11, 9),
         ap=approx_params(3, 4), rm=run_method(), act="relu6"),
    dict(kind="conv", args=dict(in_channels=16, out_channels=8, kernel_size=1, stride=1, padding=0), x=(2, 16, 7, 7),
         ap=approx_params(3, 4, dnsmp_factor=4), rm=run_method(), act=None),
    dict(kind="conv", args=dict(in_channels=8, out_channels=8, kernel_size=3, stride=1, padding=1, groups=8),
         x=(2, 8, 9, 9), ap=approx_params(3, 4), rm=run_method(), act="relu6"),          # depthwise: quirk branch
    dict(kind="conv", args=dict(in_channels=6, out_channels=6, kernel_size=3, stride=2, padding=1, groups=6),
         x=(2, 6, 9, 9), ap=approx_params(4, 3, with_s2nn2s_opt=False), rm=run_method(), act="relu6"),
    dict(kind="conv", args=dict(in_channels=8, out_channels=12, kernel_size=3, stride=1, padding=1, groups=2),
         x=(2, 8, 8, 8), ap=approx_params(4, 3), rm=run_method(), act="relu"),           # grouped, N = 6
    dict(kind="conv", args=dict(in_channels=8, out_channels=12, kernel_size=3, stride=1, padding=1), x=(2, 8, 8, 8),
         ap=approx_params(4, 3), rm=run_method(approx=False, qamaa=True), act="relu"),
    dict(kind="conv", args=dict(in_channels=4, out_channels=4, kernel_size=3, stride=1, padding=1, groups=4),
         x=(2, 4, 8, 8), ap=approx_params(4, 3), rm=run_method(approx=False, qamaa=True), act=None),  # N==1: exact
]


def qparams(ops_mod, case):
    from quantization.quantizers import FPQuantizer
    from quantization.range_estimators import RangeEstimators
    M = case["ap"]["mant_width"]
    return dict(method=FPQuantizer, act_method=FPQuantizer, n_bits=8, n_bits_act=8,
                per_channel_weights=True, weight_range_method=RangeEstimators.current_minmax.cls,
                act_range_method=RangeEstimators.allminmax.cls, quantize_input=True,
                fp8_kwargs=dict(mantissa_bits=M, set_maxval=True),
                custom_approx_params=case["ap"], run_method=case["rm"])


def build(ops_mod, case, seed):
    g = torch.Generator().manual_seed(seed)
    qp = qparams(ops_mod, case)
    if case["kind"] == "linear":
        layer = ops_mod.QCustomLinearTorch(**case["args"], bias=True, **qp)
        w = torch.randn(layer.weight.shape, generator=g) * 0.08
        b = torch.randn(layer.bias.shape, generator=g) * 0.05
        with torch.no_grad():
            layer.weight.copy_(w)
            layer.bias.copy_(b)
        extra = {}
    else:
        act = {"relu": nn.ReLU(), "relu6": nn.ReLU6(), None: None}[case.get("act")]
        layer = ops_mod.QCustomBNConv2dTorch(**case["args"], activation=act, **qp)
        w = torch.randn(layer.weight.shape, generator=g) * 0.15
        C = layer.out_channels
        extra = dict(running_mean=torch.randn(C, generator=g) * 0.1, running_var=torch.rand(C, generator=g) + 0.5,
                     gamma=torch.rand(C, generator=g) + 0.5, beta=torch.randn(C, generator=g) * 0.1)
        with torch.no_grad():
            layer.weight.copy_(w)
            layer.running_mean.copy_(extra["running_mean"])
            layer.running_var.copy_(extra["running_var"])
            layer.gamma.copy_(extra["gamma"])
            layer.beta.copy_(extra["beta"])
        b = None
    layer.eval()
    x_cal = torch.randn(case["x"], generator=g)
    x_eval = torch.randn(case["x"], generator=g) * 1.1
    if case["kind"] == "conv":   # post-ReLU-like inputs with exact zeros, as inside a network
        x_cal, x_eval = torch.relu(x_cal), torch.relu(x_eval)
    return layer, w, b, extra, x_cal, x_eval


def main():
    ops_mod = ref_shim.load_operators()
    out = {}
    meta = []
    for i, case in enumerate(CASES):
        layer, w, b, extra, x_cal, x_eval = build(ops_mod, case, 100 + i)
        sink = io.StringIO()
        with torch.inference_mode(), contextlib.redirect_stdout(sink):   # the Linear variant prints whole tensors
            layer.quantized()
            layer.estimate_ranges()
            y_cal = layer(x_cal)
            layer.fix_ranges()
            y_eval = layer(x_eval)
        out[f"w{i}"] = w.numpy()
        if b is not None:
            out[f"b{i}"] = b.numpy()
        for k, v in extra.items():
            out[f"{k}{i}"] = v.numpy()
        out[f"xcal{i}"], out[f"xeval{i}"] = x_cal.numpy(), x_eval.numpy()
        out[f"ycal{i}"], out[f"yeval{i}"] = y_cal.contiguous().numpy(), y_eval.contiguous().numpy()
        out[f"bias_a{i}"] = layer.get_acts_fp_bias().reshape(-1).numpy()
        out[f"bias_w{i}"] = layer.get_weights_fp_bias().reshape(-1).numpy()
        rb = layer.get_res_fp_bias()
        out[f"bias_r{i}"] = rb.reshape(-1).numpy() if rb is not None else np.zeros(0, np.float32)
        meta.append(dict(kind=case["kind"], args=case["args"], ap=case["ap"], rm=case["rm"], act=case.get("act")))
        print(i, case["kind"], case["args"], "rm", case["rm"], "y_eval abs max", float(y_eval.abs().max()),
              "biases", out[f"bias_a{i}"], out[f"bias_w{i}"][:4], out[f"bias_r{i}"])
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, "layers.npz"), **out)
    print("layers:", len(CASES), "cases")


if __name__ == "__main__":
    main()
