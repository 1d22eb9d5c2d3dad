"""World-size-2 (and 3) gloo tests, CPU only: the host-side sharding logic of the N>1 path
(fp8_quantization_b200/eval_harness.py) -- shard bounds, ragged all_gather of logits, range
synchronisation for sharded calibration.  No kernel is called (there is no CPU path)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n_rows, result_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from fp8_quantization_b200 import eval_harness as eh
    from fp8_quantization_b200.quantization import AllMinMaxEstimator, FPQuantizer, QuantizationManager
    torch.manual_seed(10)                       # replicated weights by seed, no broadcast
    model = torch.nn.Sequential(torch.nn.Linear(12, 7), torch.nn.ReLU(), torch.nn.Linear(7, 5))
    x = torch.randn(n_rows, 12, generator=torch.Generator().manual_seed(3))
    labels = torch.randint(0, 5, (n_rows,), generator=torch.Generator().manual_seed(4))
    logits, metrics = eh.evaluate(model, x, labels)
    with torch.inference_mode():
        want = model(x)
    ok = torch.equal(logits, want)
    top1 = float((want.argmax(1) == labels).float().mean())
    ok = ok and abs(metrics["top1"] - top1) < 1e-7
    # sharded calibration: every rank sees only its slice; merged range == range of the whole batch
    holder = torch.nn.Module()
    holder.q = QuantizationManager(qmethod=FPQuantizer, init=AllMinMaxEstimator,
                                   qparams=dict(n_bits=8, mantissa_bits=3, set_maxval=True))
    lo, hi = eh.shard_range(n_rows, rank, world)
    if hi > lo:
        holder.q.range_estimator(x[lo:hi])
    else:
        holder.q.range_estimator(x[:1] * 0)
    eh.sync_ranges(holder)
    mn, mx = torch.aminmax(x)
    ok = ok and torch.equal(holder.q.range_estimator.current_xmin, mn)
    ok = ok and torch.equal(holder.q.range_estimator.current_xmax, mx)
    ok = ok and torch.equal(holder.q.quantizer.maxval.reshape(-1), torch.max(mn.abs(), mx).reshape(-1))
    open(os.path.join(result_dir, f"r{rank}"), "w").write("ok" if ok else "bad")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_rows", [(2, 10), (2, 7), (3, 8), (2, 1)])
def test_sharded_eval_and_range_sync(tmp_path, world, n_rows):
    mp.spawn(_worker, args=(world, _free_port(), n_rows, str(tmp_path)), nprocs=world, join=True)
    assert [open(tmp_path / f"r{r}").read() for r in range(world)] == ["ok"] * world


def test_shard_range_partitions():
    from fp8_quantization_b200.eval_harness import shard_range
    for n in (0, 1, 7, 8, 256):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
