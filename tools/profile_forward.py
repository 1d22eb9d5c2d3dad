#!/usr/bin/env python
"""ONE graph replay of a calibrated model's eval forward between cudaProfilerStart/Stop, so that

  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/<name>.csv python tools/profile_forward.py --model resnet18 --batch 32

lists exactly the kernels of one forward (graph nodes are profiled one by one): the per-forward node count and the
share of each kernel.  Without ncu it prints the replay time (run it that way first).  tools/ncu_summary.py turns
the csv into the table committed under profiles/."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--mant-width", type=int, default=3)
    ap.add_argument("--no-autotune", action="store_true")
    ap.add_argument("--ofuf", action="store_true", help="opt-in v5 overflow/underflow mode (ofuf_mode = 1), BASELINE configs[3]")
    a = ap.parse_args()
    import torch
    import model_bench as mb
    import model_cases as mc
    from fp8_quantization_b200 import eval_harness as eh, ops
    dev = torch.device("cuda", 0)
    ops.set_autotune(not a.no_autotune)
    from fp8_quantization_b200 import quantization as _Q
    _Q.enable_fusions(True)      # opt-in fused forward (input quantizer, fixed-weight cache, GEMM epilogue): same bits
    M = a.mant_width
    model = mb.build(a.model, 7 - M, M, dev, ofuf=a.ofuf)
    if a.model == "bert_base":
        ids = lambda n, seed: torch.randint(0, 30522, (n, 128), generator=torch.Generator().manual_seed(seed))
        x_cal, x = ids(a.batch, 1).to(dev), ids(a.batch, 2).to(dev)
    else:
        x_cal, x = mc.images(a.batch, 224, 1).to(dev), mc.images(a.batch, 224, 2).to(dev)
    eh.calibrate(model, x_cal)
    fwd = eh.GraphedForward(model)
    with torch.inference_mode():
        for _ in range(3):
            y = fwd(x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            y = fwd(x)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        g = next(iter(fwd._graphs.values()))[0]
        torch.cuda.profiler.start()
        g.replay()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    print(json.dumps({"model": a.model, "batch": a.batch, "format": f"E{7 - M}M{M}", "ofuf_mode": int(a.ofuf), "ms_per_forward": round(ms, 3),
                      "images_per_s": round(a.batch / ms * 1e3, 1), "logits_checksum": float(y.double().sum())}))


if __name__ == "__main__":
    main()
