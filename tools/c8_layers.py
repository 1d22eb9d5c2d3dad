#!/usr/bin/env python
"""Which table variant each approx GEMM of a model's eval forward takes (byte cells or 16-bit cells), with the
escape statistics behind the device-side choice:  python tools/c8_layers.py --model vit_b16 --batch 16 --mant-width 4"""
import argparse
import collections
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="vit_b16")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--mant-width", type=int, default=4)
    a = ap.parse_args()
    import torch
    import model_bench as mb
    import model_cases as mc
    from fp8_quantization_b200 import eval_harness as eh, ops, quantization as _Q
    dev = torch.device("cuda", 0)
    _Q.enable_fusions(True)
    M = a.mant_width
    model = mb.build(a.model, 7 - M, M, dev)
    eh.calibrate(model, mc.images(a.batch, 224, 1).to(dev))
    x = mc.images(a.batch, 224, 2).to(dev)
    seen = collections.OrderedDict()
    for name in ("approx_matmul", "approx_conv2d"):
        orig = getattr(ops, name)

        def wrapped(*args, _orig=orig, _name=name, **kw):
            out = _orig(*args, **kw)
            v = ops.table_variant()
            shape = tuple(args[0].shape) + tuple(args[1].shape)
            key = (_name, shape, v["variant"], v["tmax"], v["tb_eff"])
            rec = seen.setdefault(key, {"calls": 0, "escapes": 0, "rare_b": 0})
            rec["calls"] += 1
            rec["escapes"] += v["escapes"]
            rec["rare_b"] += v["rare_b"]
            return out
        setattr(ops, name, wrapped)
    with torch.inference_mode():
        model(x)
    torch.cuda.synchronize()
    for key, rec in seen.items():
        print(json.dumps({"op": key[0], "shapes": key[1], "variant": key[2], "tmax": key[3], "tb_eff": key[4], **rec}))


if __name__ == "__main__":
    main()
