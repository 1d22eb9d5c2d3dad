"""N > 1 on real GPUs (run with `-m gpu` on a box with at least 2 B200s; skipped otherwise): two NCCL ranks, one
per GPU, evaluate disjoint halves of a batch through the CUDA-graph runner with the logits all_gather captured in
the graph (eval_harness.GraphedForward), and the gathered logits must be BIT-IDENTICAL to a single-GPU forward of
the whole batch -- rows are independent given replicated weights and ranges."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, result_dir):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    import model_cases as mc
    import test_gpu_models as tgm
    from fp8_quantization_b200 import eval_harness as eh
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    res, per = 64, 4
    model = tgm.build("resnet18", res, 4, 3).to(dev)          # replicated by seed
    eh.calibrate(model, mc.images(per, res, 1).to(dev))          # replicated calibration batch
    n = per * world
    x = mc.images(n, res, 2).to(dev)
    ok = True
    fwd = eh.GraphedForward(model, gather_total=n)
    for forward in (None, fwd, fwd):                             # eager gather; graph capture; graph replay
        logits, _ = eh.evaluate(model, x, forward=forward)
        with torch.inference_mode():
            whole = model(x)
        ok = ok and logits.shape == whole.shape and torch.equal(logits, whole)
    open(os.path.join(result_dir, f"r{rank}"), "w").write(f"{'ok' if ok else 'bad'} {fwd.collective_in_graph}")
    dist.barrier()
    # a live CUDA graph that holds captured NCCL kernels keeps the communicator busy: destroy_process_group() then
    # waits for ever -- drop the graphs first
    import gc
    fwd._graphs.clear()
    del fwd, logits
    gc.collect()
    torch.cuda.synchronize()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_two_rank_gathered_logits_equal_single_gpu(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = [open(tmp_path / f"r{r}").read().split() for r in range(world)]
    assert [g[0] for g in got] == ["ok"] * world, got
    print("all_gather captured inside the CUDA graph:", got[0][1])
