#!/usr/bin/env python
"""Model-level throughput of the approx-sim eval forward (images/s) on synthetic ImageNet-shaped data.

  python tools/model_bench.py --model resnet18 --batch 8 [--graph] [--steps 10]
Used by bench.py for the auxiliary `models` block; also a standalone profiling target."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MACS_PER_IMAGE = {"resnet18": 1.814e9, "mobilenetv2": 0.301e9, "vit_b16": 16.73e9, "bert_base": 10.87e9}   # bert_base: per sequence of 128 tokens, NON-approx (tcgen05)


def build(model_name, E, M, device, ofuf=False):
    """ofuf: the opt-in v5 overflow / underflow mode (BASELINE configs[3] "with OF/UF sim"):
    sim_hw_add_OFUF / with_OF_opt / with_UF_opt on and LIVE (ofuf_mode = 1)."""
    import torch
    import model_cases as mc
    from fp8_quantization_b200 import model_wrap as mw
    cap = mc.approx_params(E, M)
    if ofuf:
        cap.update(sim_hw_add_OFUF=True, with_OF_opt=True, with_UF_opt=True, ofuf_mode=1)
    qp = dict(method="fp_quantizer", act_method="fp_quantizer", n_bits=8, n_bits_act=8, per_channel_weights=True,
              weight_range_method="current_minmax", act_range_method="allminmax", quantize_input=True,
              fp8_kwargs=dict(mantissa_bits=M, set_maxval=True), custom_approx_params=cap,
              run_method=dict(mc.RUN_APPROX))
    if model_name == "resnet18":
        model = mw.QuantizedResNet(mc.fp_resnet18(), **qp)
    elif model_name == "bert_base":
        from fp8_quantization_b200 import bert
        torch.manual_seed(mc.SEED)
        qp["run_method"] = dict(bert.RUN_PLAIN)
        model = bert.bert_quantized(**qp)
    elif model_name == "vit_b16":
        from fp8_quantization_b200 import vit
        torch.manual_seed(mc.SEED)
        model = vit.vit_quantized_approx(**qp)
    else:
        model = mw.QuantizedMobileNetV2(mc.fp_mobilenetv2(mw.MobileNetV2, 224), **qp)
    return model.to(device).eval()


def images_per_s(model_name="resnet18", batch=8, steps=10, warmup=3, E=4, M=3, graph=False, device=None,
                 calib_batch=None):
    import torch
    import model_cases as mc
    from fp8_quantization_b200 import eval_harness as eh, ops
    device = device or torch.device("cuda", torch.cuda.current_device())
    ops.set_autotune(True)
    from fp8_quantization_b200 import quantization as _Q
    _Q.enable_fusions(True)      # opt-in fused forward (input quantizer, fixed-weight cache, GEMM epilogue): same bits
    model = build(model_name, E, M, device)
    if model_name == "bert_base":
        ids = lambda n, seed: torch.randint(0, 30522, (n, 128), generator=torch.Generator().manual_seed(seed))
        x_cal, x = ids(calib_batch or batch, 1).to(device), ids(batch, 2).to(device)
    else:
        x_cal = mc.images(calib_batch or batch, 224, 1).to(device)
        x = mc.images(batch, 224, 2).to(device)
    eh.calibrate(model, x_cal)
    with torch.inference_mode():
        for _ in range(warmup):
            y = model(x)
        torch.cuda.synchronize()
        run = lambda: model(x)
        if graph:
            g = torch.cuda.CUDAGraph()
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                model(x)
            torch.cuda.current_stream().wait_stream(s)
            with torch.cuda.graph(g):
                y = model(x)
            run = g.replay
            run()
            torch.cuda.synchronize()
        l0 = ops.kernel_launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            run()
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1) / steps
    return {"model": model_name, "batch": batch, "ms_per_forward": round(ms, 3), "images_per_s": round(batch / ms * 1e3, 1),
            "approx_gmacs": round(batch * MACS_PER_IMAGE[model_name] / ms / 1e6, 1), "cuda_graph": bool(graph),
            "kernel_launches_per_forward": (ops.kernel_launches() - l0) // max(steps, 1) if not graph else None,
            "wall_ms_per_forward": round(wall / steps * 1e3, 3), "format": f"E{E}M{M}",
            "logits_checksum": float(y.double().sum())}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18", choices=list(MACS_PER_IMAGE))
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mant-width", type=int, default=3)
    ap.add_argument("--graph", action="store_true")
    a = ap.parse_args()
    print(json.dumps(images_per_s(a.model, a.batch, a.steps, a.warmup, 7 - a.mant_width, a.mant_width, a.graph)))
