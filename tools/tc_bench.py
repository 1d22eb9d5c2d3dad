#!/usr/bin/env python
"""Throughput of the non-approx tensor-core matmul (K3, csrc/fq_gemm_tc.cu) on fake-quantized operands.

  python tools/tc_bench.py [--shapes 4096x4096x4096,25216x768x3072] [--steps 10]
Reports, per shape: whole call (bf16 convert pass + tcgen05 GEMM) and torch's fp32 `x @ y` (what the
reference runs, cuBLAS) in ms and TFLOP/s, and the fraction of the measured dense bf16 peak
(MEASURED_PEAKS.json)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="4096x4096x4096,25216x768x3072,25216x3072x768,32768x768x768")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--mant-width", type=int, default=3)
    a = ap.parse_args()
    import torch
    from fp8_quantization_b200 import ops
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
    except Exception:
        pass
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(10)
    mant = torch.tensor([float(a.mant_width)])
    for shp in a.shapes.split(","):
        Mr, K, N = (int(v) for v in shp.split("x"))
        x = (torch.randn(Mr, K, generator=g) * 0.5).to(dev)
        w = (torch.randn(N, K, generator=g) * 0.05).to(dev)
        xq, _ = ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
        wq, _ = ops.fp8_fakequant(w, 8, w.abs().amax(dim=1), mant, 1)
        out = torch.empty((Mr, N), device=dev)

        def timed(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / a.steps

        ms = timed(lambda: ops.fq_matmul(xq, wq.t(), out=out))
        path = ops.last_path()
        ms_t = timed(lambda: torch.matmul(xq, wq.t(), out=out))
        fl = 2.0 * Mr * K * N
        err = float((ops.fq_matmul(xq, wq.t()).double() - xq.double() @ wq.double().t()).abs().max())
        print(json.dumps({"shape": shp, "path": path, "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1),
                          "frac_of_measured_bf16_peak": round(fl / ms / 1e9 / peak, 3) if peak else None,
                          "torch_fp32_ms": round(ms_t, 4), "torch_fp32_tflops": round(fl / ms_t / 1e9, 1),
                          "speedup_vs_torch_fp32": round(ms_t / ms, 2), "max_abs_err_vs_fp64": err}), flush=True)


if __name__ == "__main__":
    main()
