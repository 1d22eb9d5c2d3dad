#!/usr/bin/env python
"""HBM roofline of the FP8 fake-quantizer kernel (K4, SURVEY.md 8(d): 8 B per element) and of the fused
bn_act epilogue: achieved GB/s against MEASURED_PEAKS.json's copy bandwidth.

  python tools/fakequant_bench.py [--elements 268435456]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--elements", type=int, default=1 << 28)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    import torch
    from fp8_quantization_b200 import ops
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6550.0
    x = torch.randn(a.elements, device="cuda")
    mv = x.abs().max().reshape(1)
    out = {}
    for name, mant, fn in (("fakequant_tensor_E4M3", 3.0, None), ("fakequant_tensor_E2M5", 5.0, None)):
        m = torch.tensor([mant])
        for _ in range(3):
            ops.fp8_fakequant(x, 8, mv, m, 1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            ops.fp8_fakequant(x, 8, mv, m, 1)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        gbs = 8.0 * a.elements / ms / 1e6
        out[name] = {"ms": round(ms, 4), "GB/s": round(gbs, 1), "frac_of_measured_copy_peak": round(gbs / peak, 4)}
    print(json.dumps({"elements": a.elements, "alg_bytes_per_element": 8, "peak_GB/s": peak, **out}))


if __name__ == "__main__":
    main()
