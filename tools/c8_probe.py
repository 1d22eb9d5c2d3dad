#!/usr/bin/env python
"""One approx GEMM per variant (default dispatch, then params.force_table16) on bench.py-style operands, for an ncu
launch list:  ncu --metrics gpu__time_duration.sum --csv python tools/c8_probe.py --shape 3152 768 768 --mant 4"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from fp8_quantization_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--shape", type=int, nargs=3, default=[3152, 768, 768])
ap.add_argument("--mant", type=int, default=4)
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
rows, K, N = a.shape
M = a.mant
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(10)
x = (torch.randn(rows, K, generator=g) * 0.5).to(dev)
w = (torch.randn(N, K, generator=g) * 0.02).to(dev)
mant = torch.tensor([float(M)])
xq, bA = ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
wq, bB = ops.fp8_fakequant(w, 8, w.abs().amax(dim=1), mant, 1)
mv = ((xq[:256] @ wq.t()).abs().max() * 1.5).reshape(1)
_, bR = ops.fp8_fakequant(mv.clone(), 8, mv, mant, 1)
out = torch.empty((rows, N), device=dev)
for force in (False, True):
    p = ops.make_params(bench.approx_params(M), force_table16=force)
    for _ in range(a.reps):
        ops.approx_matmul(xq, wq.t(), bA, bB, bR, p, out=out)
    torch.cuda.synchronize()
    print("force_table16", force, ops.last_path(), ops.table_variant(), float(out.double().sum()))
