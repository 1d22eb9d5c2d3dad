#!/usr/bin/env python
"""Per-shape timing of the approx GEMM over the GEMM shapes of a model.

  python tools/layer_bench.py --model resnet18 --batch 32 [--zero-frac 0.5] [--mant-width 3]
For every distinct (Mrows, K, N): whole call (predecode + table GEMM) and the table GEMM alone
(CUDA events on the launching stream), in ms and T approx-MAC/s.  Operands are sampled on the FP8
grid; `--zero-frac` of the activations are zero (post-ReLU tensors are about half zeros)."""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# (count, Mrows per image, K, N)
SHAPES = {
    "resnet18": [(1, 12544, 147, 64), (4, 3136, 576, 64), (1, 784, 576, 128), (3, 784, 1152, 128),
                 (1, 196, 1152, 256), (3, 196, 2304, 256), (1, 49, 2304, 512), (3, 49, 4608, 512),
                 (1, 784, 64, 128), (1, 196, 128, 256), (1, 49, 256, 512)],
    "vit_b16": [(48, 197, 768, 768), (12, 197, 768, 3072), (12, 197, 3072, 768)],
    "mobilenetv2_pw": [(1, 12544, 32, 16), (1, 12544, 16, 96), (1, 3136, 96, 24), (2, 3136, 24, 144),
                       (1, 3136, 144, 24), (1, 784, 144, 32), (3, 784, 32, 192), (2, 784, 192, 32),
                       (1, 196, 192, 64), (4, 196, 64, 384), (3, 196, 384, 64), (1, 196, 384, 96),
                       (3, 196, 96, 576), (2, 196, 576, 96), (1, 49, 576, 160), (3, 49, 160, 960),
                       (2, 49, 960, 160), (1, 49, 960, 320), (1, 49, 320, 1280)],
    "square": [(1, 4096, 4096, 4096)],
}

# conv form (--conv): (count, C_in, H, C_out, k, stride, pad)
CONVS = {
    "resnet18": [(1, 3, 224, 64, 7, 2, 3), (4, 64, 56, 64, 3, 1, 1), (1, 64, 56, 128, 3, 2, 1), (3, 128, 28, 128, 3, 1, 1),
                 (1, 128, 28, 256, 3, 2, 1), (3, 256, 14, 256, 3, 1, 1), (1, 256, 14, 512, 3, 2, 1), (3, 512, 7, 512, 3, 1, 1),
                 (1, 64, 56, 128, 1, 2, 0), (1, 128, 28, 256, 1, 2, 0), (1, 256, 14, 512, 1, 2, 0)],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18", choices=sorted(SHAPES))
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--mant-width", type=int, default=3)
    ap.add_argument("--zero-frac", type=float, default=0.5)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--mode", default="approx", choices=["approx", "qamaa"])
    ap.add_argument("--check", action="store_true", help="compare with the arithmetic kernel, bit for bit")
    ap.add_argument("--conv", action="store_true", help="conv form (implicit im2col) over the model's conv geometries")
    a = ap.parse_args()
    import torch
    import fp8_quantization_b200 as pkg
    from fp8_quantization_b200 import ops
    lib = pkg._lib.load()
    dev = torch.device("cuda", 0)
    M = a.mant_width
    E = 7 - M
    # mant_width 2 (E5M2) has no reference answer (approx_v9 raises for it): all-zero table, "parity unpinned"
    params = ops.make_params(dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True,
                                  with_s2nn2s_opt=True, quant_btw_mult_accu=True), allow_unpinned_format=(M == 2))
    g = torch.Generator(device="cpu").manual_seed(10)
    mant = torch.tensor([float(M)])
    tot_ms = tot_kern = tot_macs = 0.0
    shapes = SHAPES[a.model] if not a.conv else CONVS[a.model]
    for shape in shapes:
        if a.conv:
            cnt, Ci, H, N, k, st, pad = shape
            Ho = (H + 2 * pad - k) // st + 1
            rows, K = a.batch * Ho * Ho, Ci * k * k
            x = torch.randn(a.batch, Ci, H, H, generator=g).abs_() * 0.5
            x[torch.rand(x.shape, generator=g) < a.zero_frac] = 0.0
            w = torch.randn(N, Ci, k, k, generator=g) * 0.05
            x, w = x.to(dev), w.to(dev)
            xq, bA = ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
            wq, bB = ops.fp8_fakequant(w.reshape(N, -1), 8, w.reshape(N, -1).abs().amax(dim=1), mant, 1)
            wq = wq.reshape(w.shape)
            ref = torch.nn.functional.conv2d(xq[:1], wq, stride=st, padding=pad)
            _, bR = ops.fp8_fakequant(ref, 8, (ref.abs().max() * 2).reshape(1), mant, 1)
            outs = []
            step = lambda: outs.__setitem__(slice(None), [ops.approx_conv2d(xq, wq, bA, bB, bR, params, (st, st), (pad, pad), (1, 1), 1)])
            if a.mode == "qamaa":
                mv = (ref.abs().max() * 2).reshape(1)
                step = lambda: outs.__setitem__(slice(None), [ops.qamaa_conv2d(xq, wq, 8, mv, mant, 1, (st, st), (pad, pad), (1, 1), 1, x_bias=bA, w_bias=bB)])
        else:
            cnt, mr, K, N = shape
            rows = mr * (a.batch if a.model != "square" else 1)
            x = torch.randn(rows, K, generator=g).abs_() * 0.5
            x[torch.rand(rows, K, generator=g) < a.zero_frac] = 0.0
            w = torch.randn(N, K, generator=g) * 0.05
            x, w = x.to(dev), w.to(dev)
            xq, bA = ops.fp8_fakequant(x, 8, x.abs().max().reshape(1), mant, 1)
            wq, bB = ops.fp8_fakequant(w, 8, w.abs().amax(dim=1), mant, 1)
            ref = xq[:64] @ wq.t()
            _, bR = ops.fp8_fakequant(ref, 8, (ref.abs().max() * 2).reshape(1), mant, 1)
            out = torch.empty((rows, N), device=dev)
            step = lambda: ops.approx_matmul(xq, wq.t(), bA, bB, bR, params, out=out)
            if a.mode == "qamaa":
                mv = (ref.abs().max() * 2).reshape(1)
                step = lambda: ops.qamaa_matmul(xq, wq.t(), 8, mv, mant, 1, out=out, x_bias=bA, y_bias=bB)
        for _ in range(2):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        lib.b200_profile_begin()
        for _ in range(a.steps):
            step()
        t, c = ctypes.c_double(0), ctypes.c_int(0)
        lib.b200_profile_end(ctypes.byref(t), ctypes.byref(c))
        kms = t.value / max(c.value, 1)
        macs = rows * K * N
        if a.check and not a.conv:
            pg = ops.make_params(dict(expo_width=E, mant_width=M, dnsmp_factor=3, withComp=True, with_approx=True,
                                      with_s2nn2s_opt=True, quant_btw_mult_accu=True), force_generic=True)
            want = ops.approx_matmul(xq, wq.t(), bA, bB, bR, pg)
            nbad = int((want.view(torch.int32) != out.view(torch.int32)).sum())
            print("check", rows, K, N, "mismatching elements:", nbad, "max abs diff", float((want - out).abs().max()))
        print(json.dumps({"rows": rows, "K": K, "N": N, "count": cnt, "path": ops.last_path(), "ms": round(ms, 4),
                          "gemm_ms": round(kms, 4), "tmacs": round(macs / ms / 1e9, 3),
                          "gemm_tmacs": round(macs / kms / 1e9, 3) if kms else None}), flush=True)
        tot_ms += cnt * ms
        tot_kern += cnt * kms
        tot_macs += cnt * macs
    print(json.dumps({"model": a.model, "batch": a.batch, "format": f"E{E}M{M}", "zero_frac": a.zero_frac,
                      "sum_ms": round(tot_ms, 3), "sum_gemm_ms": round(tot_kern, 3),
                      "tmacs": round(tot_macs / tot_ms / 1e9, 3)}))


if __name__ == "__main__":
    main()
