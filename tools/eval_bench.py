#!/usr/bin/env python
"""Model evaluation throughput over N GPUs of one box (BASELINE.json metric, second half:
"ResNet-18 approx-sim images/sec at 1/2/4/8 B200").

  python tools/eval_bench.py --model resnet18 --per-gpu-batch 32
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/eval_bench.py --model resnet18 --per-gpu-batch 32

One process per GPU.  Weights are replicated by seed, the calibration batch is replicated (bit-identical
ranges on every rank), the eval batch of N * per_gpu_batch synthetic images is sharded by rows, and the
only data-path collective is ONE all_gather of the logits per step (NCCL over NVLink; latency bound).
Weak scaling: per-GPU work is fixed.  Timing: CUDA events around K steps, barrier + synchronize on both
sides, max over ranks.  Rank 0 also checks that the gathered logits equal a single-GPU forward of the
whole batch (rows are independent: the sharded result must be bit-identical)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--per-gpu-batch", type=int, default=32)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mant-width", type=int, default=3)
    ap.add_argument("--check", action="store_true", help="rank 0 re-runs the whole batch alone and compares")
    ap.add_argument("--graph", action="store_true", help="replay the forward as a CUDA graph (eval_harness.GraphedForward)")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    import model_bench as mb
    import model_cases as mc
    from fp8_quantization_b200 import eval_harness as eh, ops
    ops.set_autotune(True)
    from fp8_quantization_b200 import quantization as _Q
    _Q.enable_fusions(True)      # opt-in fused forward (input quantizer, fixed-weight cache, GEMM epilogue): same bits

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    M = a.mant_width
    model = mb.build(a.model, 7 - M, M, dev)
    n = a.per_gpu_batch * world
    if a.model == "bert_base":
        mk = lambda k, seed: torch.randint(0, 30522, (k, 128), generator=torch.Generator().manual_seed(seed))
    else:
        mk = lambda k, seed: mc.images(k, 224, seed)
    x_cal = mk(a.per_gpu_batch, 1).to(dev)          # replicated calibration batch
    x = mk(n, 2).to(dev)                            # every rank holds the batch, evaluates its slice
    eh.calibrate(model, x_cal)
    fwd = eh.GraphedForward(model, gather_total=n if world > 1 else None) if a.graph else None
    for _ in range(max(a.warmup, 3)):
        logits, _ = eh.evaluate(model, x, forward=fwd)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        logits, _ = eh.evaluate(model, x, forward=fwd)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_step = float(ms.item()) / a.steps
    same = None
    if a.check and rank == 0 and world > 1:
        with torch.inference_mode():
            whole = model(x)
        same = bool(torch.equal(whole, logits))      # `logits`: the last timed step's gathered result
    if rank == 0:
        print(json.dumps({"metric": f"{a.model}_approx_sim_images_per_s", "value": round(n / ms_step * 1e3, 1),
                          "unit": "images/s", "n_gpus": world, "steps": a.steps, "ms_per_step": round(ms_step, 3),
                          "scaling": "weak", "cuda_graph": bool(a.graph), "per_gpu_batch": a.per_gpu_batch, "global_batch": n,
                          "format": f"E{7 - M}M{M}", "approx_gmacs": round(n * mb.MACS_PER_IMAGE[a.model] / ms_step / 1e6, 1),
                          "collective": ("one all_gather of logits per step" + (", captured in the CUDA graph" if fwd is not None and fwd.collective_in_graph else "")) if world > 1 else None,
                          "sharded_equals_single_gpu": same, "logits_checksum": float(logits.double().sum())}),
              flush=True)
    if world > 1:
        if fwd is not None:      # a live graph with captured NCCL kernels makes destroy_process_group() wait for ever
            fwd._graphs.clear()
        del fwd
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
