#!/usr/bin/env python
"""Before / after of the byte-cell table variant (csrc/lut_gemm.cuh, C8): whole step (operand predecode + table GEMM)
of bench.py's sweep operands and of ViT-B/16-shaped GEMMs, default dispatch against params.force_table16.
  python tools/c8_bench.py [--steps 5] [--zeros 0.0]"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fp8_quantization_b200 import ops  # noqa: E402


def timed(step, steps):
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--zeros", type=float, default=0.0, help="fraction of activations set to zero (post-ReLU tensors)")
    ap.add_argument("--autotune", type=int, default=1)
    ap.add_argument("--mants", type=int, nargs="+", default=[4, 5])
    ap.add_argument("--shape", type=int, nargs=3, action="append", help="rows K N (repeatable); default: 4096^3 and ViT-B shapes")
    a = ap.parse_args()
    ops.set_autotune(bool(a.autotune))
    dev = torch.device("cuda", 0)
    shapes = [tuple(t) for t in a.shape] if a.shape else [(4096, 4096, 4096), (3152, 768, 768), (3152, 768, 3072), (3152, 3072, 768)]
    for M in a.mants:
        for (rows, K, N) in shapes:
            g = torch.Generator().manual_seed(10)
            x = torch.randn(rows, K, generator=g) * 0.5
            if a.zeros > 0:
                x = x * (torch.rand(rows, K, generator=g) >= a.zeros)
            w = torch.randn(N, K, generator=g) * 0.02
            x, w = x.to(dev), w.to(dev)
            mant = torch.tensor([float(M)])
            xq, bA = ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
            wq, bB = ops.fp8_fakequant(w, 8, w.abs().amax(dim=1), mant, 1)
            mv = ((xq[:256] @ wq.t()).abs().max() * 1.5).reshape(1)
            _, bR = ops.fp8_fakequant(mv.clone(), 8, mv, mant, 1)
            out = torch.empty((rows, N), device=dev)
            rec = {"mant_width": M, "shape": [rows, K, N], "zeros": a.zeros}
            for name, force in (("byte_cells", False), ("cells16", True)):
                p = ops.make_params(bench.approx_params(M), force_table16=force)
                ms = timed(lambda: ops.approx_matmul(xq, wq.t(), bA, bB, bR, p, out=out), a.steps)
                rec[name] = {"ms": round(ms, 4), "gmacs": round(rows * K * N / ms / 1e6, 1), "path": ops.last_path(),
                             **{k: v for k, v in ops.table_variant().items() if k in ("variant", "escapes", "tmax")}}
                rec[name + "_sum"] = float(out.double().sum())
            ms = timed(lambda: ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1, out=out, x_bias=bA, y_bias=bB), a.steps)
            rec["qamaa"] = {"ms": round(ms, 4), "gmacs": round(rows * K * N / ms / 1e6, 1), "path": ops.last_path(),
                            "variant": ops.table_variant()["variant"]}
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
