"""Synthetic-batch evaluation of a quantized model, sharded by batch across ranks.

The reference is single-process (image_net.py:59-202: estimate_ranges -> one calibration batch ->
fix_ranges -> eval batches).  Rows (images) are independent given replicated weights and calibrated
ranges, so N GPUs run N replicas on disjoint slices of the eval batch:
  * weights are replicated by seeding every rank identically (no broadcast);
  * calibration runs on the SAME full calibration batch on every rank (replicated: bit-identical
    ranges, one extra forward) -- or, with `shard_calibration=True`, on each rank's slice followed by
    an all_reduce(MIN/MAX) of every range estimator before fix_ranges();
  * the only data-path collective is ONE all_gather of the logits (latency bound, nothing to fuse).
Works with any torch.distributed backend (NCCL on the B200 box, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist

from .quantization import QuantizationManager


def shard_range(n, rank, world):
    """Contiguous, balanced slice [lo, hi) of n items for `rank` (first n % world ranks get one more)."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def sync_ranges(model, group=None):
    """all_reduce(MIN) / all_reduce(MAX) of every range estimator's running min / max, then push the
    merged range into its quantizer (quantization/range_estimators.py:56-100 state)."""
    rank, world = _world(group)
    if world == 1:
        return
    for m in model.modules():
        if isinstance(m, QuantizationManager) and m.range_estimator is not None \
                and m.range_estimator.current_xmin is not None:
            est = m.range_estimator
            dist.all_reduce(est.current_xmin, op=dist.ReduceOp.MIN, group=group)
            dist.all_reduce(est.current_xmax, op=dist.ReduceOp.MAX, group=group)
            m.set_quant_range(est.current_xmin, est.current_xmax)


def calibrate(model, calib_batch, shard_calibration=False, group=None):
    """estimate_ranges -> forward -> fix_ranges (image_net.py:100-112)."""
    rank, world = _world(group)
    model.quantized()
    model.estimate_ranges()
    x = calib_batch
    if shard_calibration and world > 1:
        lo, hi = shard_range(x.shape[0], rank, world)
        x = x[lo:hi]
    with torch.inference_mode():
        out = model(x)
    if shard_calibration:
        sync_ranges(model, group)
    model.fix_ranges()
    return out


_GATHER_BUFS = {}


def gather_rows(local, n_total, group=None):
    """all_gather of per-rank row blocks (pads to the largest block when the shards are ragged).  The receive
    buffer is allocated once per (shape, dtype, device) and reused: with equal shards the result is a VIEW of it
    (valid until the next call with the same geometry) and no padding or concatenation pass runs."""
    rank, world = _world(group)
    if world == 1:
        return local
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    width = max(hi - lo for lo, hi in sizes)
    tail = tuple(local.shape[1:])
    key = (world, width, tail, local.dtype, local.device)
    out = _GATHER_BUFS.get(key)
    if out is None:
        out = _GATHER_BUFS[key] = local.new_empty((world * width,) + tail)
    if all(hi - lo == width for lo, hi in sizes):
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    padded = local.new_zeros((width,) + tail)
    padded[: local.shape[0]] = local
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * width: r * width + (hi - lo)] for r, (lo, hi) in enumerate(sizes)], dim=0)


class GraphedForward:
    """The eval forward of a calibrated model (fixed ranges) captured in a CUDA graph, one graph per input
    shape.  A ResNet-18 forward at batch 8 is ~200 short launches: replaying them as one graph removes the
    host launch cost between kernels.  Only for models whose forward is launch-only on fixed ranges (no host
    reads): every operator of this package qualifies (device-side dispatch flags, biases read on the
    device).  The captured kernels use the package's shared workspace; the runner keeps it alive.

    gather_total: with an initialised process group and equal shards, the ONE data-path collective -- the
    all_gather of the logits -- is captured INSIDE the graph, after the forward, into a
    pre-allocated buffer: a step is then a single graph launch.  The call returns the gathered [n_total, classes]
    logits.  Falls back to an eager all_gather after the replay if the collective cannot be captured.

    The returned tensor is a VIEW of the graph's static output: consume (or clone) it before the next call."""

    def __init__(self, model, gather_total=None, group=None):
        self.model = model
        self.gather_total = gather_total
        self.group = group
        self.collective_in_graph = None
        self._graphs = {}

    def __call__(self, x):
        key = (tuple(x.shape), x.dtype, x.device)
        ent = self._graphs.get(key)
        if ent is None:
            ent = self._capture(x)
            self._graphs[key] = ent
        g, x_static, y_static, gathered, _ = ent
        x_static.copy_(x)
        g.replay()
        if gathered is not None:
            return gathered
        if self.gather_total is not None:
            return gather_rows(y_static.float(), self.gather_total, self.group)
        return y_static

    def _capture(self, x):
        from . import ops
        x_static = x.clone()
        side = torch.cuda.Stream(device=x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side), torch.inference_mode():   # warm-up: table builds, workspace growth
            self.model(x_static)
            y = self.model(x_static)
            rank, world = _world(self.group)
            want_gather = self.gather_total is not None and world > 1 and \
                self.gather_total == world * x.shape[0] and y.dim() == 2
            if want_gather:
                gathered = y.new_empty((self.gather_total, y.shape[1]), dtype=torch.float32)
                dist.all_gather_into_tensor(gathered, y.float().contiguous(), group=self.group)   # communicator warm-up
        torch.cuda.current_stream(x.device).wait_stream(side)
        if want_gather:
            try:
                g = torch.cuda.CUDAGraph()
                with torch.inference_mode(), torch.cuda.graph(g):
                    y_static = self.model(x_static)
                    dist.all_gather_into_tensor(gathered, y_static.float().contiguous(), group=self.group)
                self.collective_in_graph = True
                return g, x_static, y_static, gathered, self._keepalive(ops)
            except Exception:    # the collective is not capturable on this stack: graph the forward only
                self.collective_in_graph = False
                torch.cuda.synchronize(x.device)
        g = torch.cuda.CUDAGraph()
        with torch.inference_mode(), torch.cuda.graph(g):
            y_static = self.model(x_static)
        return g, x_static, y_static, None, self._keepalive(ops)

    def _keepalive(self, ops):
        """What the captured kernels read by address and nobody else is guaranteed to keep: the package's workspaces and
        the layers' cached quantized weights (quantization.CACHE_FIXED_WEIGHTS; a later forward with another setting, or
        a weight update, drops the cache entry -- the graph must still own the tensors it was captured with)."""
        cached = [m.__dict__["_b200_wcache"] for m in self.model.modules() if "_b200_wcache" in m.__dict__]
        # ... and the quantizers' range / exponent-bias tensors: plain attributes that a later eager forward REPLACES
        # (FPQuantizer.forward assigns a fresh custom_bias on every call, fp8_quantizer.py:244)
        for m in self.model.modules():
            cached += [t for t in (getattr(m, "custom_bias", None), getattr(m, "maxval", None)) if torch.is_tensor(t)]
        # ... and the decoded weight planes ops keeps per layer (approx_matmul / approx_conv2d `prepared`)
        return dict(ops._WORKSPACES), cached, dict(ops._BPLANES), {k: v[1] for k, v in ops._PREPARED.items()}


_N_CLASSES = {}


def evaluate(model, images, labels=None, group=None, forward=None):
    """Forward this rank's slice of `images`, gather the logits of the whole batch on every rank.
    Returns (logits [B, classes], metrics dict with top-1 / top-5 when labels are given).
    `forward`: a callable used instead of `model` (e.g. a GraphedForward of it; one built with
    gather_total=len(images) does the gather inside its graph).
    The class count an empty shard needs for the collective is agreed on ONCE per model (one all_reduce in the
    first call), not every step; with equal shards the result is a view of a reused receive buffer."""
    rank, world = _world(group)
    n = images.shape[0]
    lo, hi = shard_range(n, rank, world)
    fwd = forward or model
    with torch.inference_mode():
        local = fwd(images[lo:hi]) if hi > lo else None
    if isinstance(fwd, GraphedForward) and fwd.gather_total == n and world > 1 and local is not None \
            and local.shape[0] == n:
        logits = local                                   # gathered inside the graph (or right after the replay)
    else:
        if world > 1:
            ncls = _N_CLASSES.get(id(model))
            if ncls is None:
                cls = torch.tensor([local.shape[1] if local is not None else 0], dtype=torch.int64).to(images.device)
                dist.all_reduce(cls, op=dist.ReduceOp.MAX, group=group)
                ncls = _N_CLASSES[id(model)] = int(cls.item())
            if local is None:
                local = images.new_zeros((0, ncls))
        elif local is None:
            local = images.new_zeros((0, 1))
        logits = gather_rows(local.float().contiguous(), n, group)
    metrics = {}
    if labels is not None:
        top5 = logits.topk(min(5, logits.shape[1]), dim=1).indices
        metrics["top1"] = float((top5[:, 0] == labels).float().mean())
        metrics["top5"] = float((top5 == labels[:, None]).any(dim=1).float().mean())
    return logits, metrics
