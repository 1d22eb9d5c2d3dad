"""Import shim for the UNMODIFIED reference (revollllt/FP8_quantization).

Test infrastructure only (like everything under oracle/).  Users:
  * tests/golden/make_*.py       generate the committed golden vectors from /root/reference (build container);
  * tests/test_gpu_dropin.py     the drop-in test on the reference's own classes (GPU box: oracle/_ref);
  * bench.py                     the cpu_baseline / gpu_eager_reference legs and `--impl reference`.
The product package (fp8_quantization_b200/) never imports this file.

Where the reference comes from: $FP8_REFERENCE_ROOT, else /root/reference (build container), else
oracle/_ref/ -- the git-ignored copy tools/install_ref.sh makes, which ships with the gpurun snapshot.

Blockers worked around:
  1. ``approx/approx_calculation.py:6``  ``import cupy``            -> empty stub module
  2. ``quantization/hijacker.py:7-8``    ``timm`` activation classes -> six empty nn.Module stubs
  3. ``approx/approx_matmul_whole_v9.py:393-549,560-589`` build tables with ``device='cuda'``
     -> ``torch.tensor`` / ``torch.zeros`` are wrapped to drop ``device='cuda'`` when CUDA is absent.
On a box with a GPU the reference runs unmodified on 'cuda' (only stubs 1-2 apply); a process that
wants the reference's CPU path there hides the GPU first (CUDA_VISIBLE_DEVICES="" before importing torch).
"""
import os
import sys
import types

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))


def _find_root():
    for cand in (os.environ.get("FP8_REFERENCE_ROOT"), "/root/reference", os.path.join(_HERE, "_ref")):
        if cand and os.path.isdir(os.path.join(cand, "approx")):
            return cand
    return os.path.join(_HERE, "_ref")


REF_ROOT = _find_root()


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "approx"))


def _install_stubs():
    if "cupy" not in sys.modules:
        import importlib.machinery
        sys.modules["cupy"] = types.ModuleType("cupy")
        sys.modules["cupy"].__spec__ = importlib.machinery.ModuleSpec("cupy", loader=None)
    if "timm" not in sys.modules:
        timm = types.ModuleType("timm")
        models = types.ModuleType("timm.models")
        layers = types.ModuleType("timm.models.layers")
        act = types.ModuleType("timm.models.layers.activations")
        act_me = types.ModuleType("timm.models.layers.activations_me")
        for name in ("Swish", "HardSwish", "HardSigmoid"):
            setattr(act, name, type(name, (torch.nn.Module,), {}))
        for name in ("SwishMe", "HardSwishMe", "HardSigmoidMe"):
            setattr(act_me, name, type(name, (torch.nn.Module,), {}))
        timm.models = models
        models.layers = layers
        layers.activations = act
        layers.activations_me = act_me
        stubs = {
            "timm": timm, "timm.models": models, "timm.models.layers": layers,
            "timm.models.layers.activations": act, "timm.models.layers.activations_me": act_me,
        }
        import importlib.machinery
        for name, mod in stubs.items():     # a module without a spec makes importlib.util.find_spec raise (transformers probes timm)
            mod.__spec__ = importlib.machinery.ModuleSpec(name, loader=None)
        sys.modules.update(stubs)


_patched = False


def _drop_cuda_device():
    """Make torch.tensor/zeros ignore device='cuda' when no GPU exists (tables are built on 'cuda')."""
    global _patched
    if _patched or torch.cuda.is_available():
        return
    _patched = True
    for fname in ("tensor", "zeros"):
        orig = getattr(torch, fname)

        def wrapped(*a, __orig=orig, **kw):
            if str(kw.get("device", "")) == "cuda":
                kw.pop("device")
            return __orig(*a, **kw)

        setattr(torch, fname, wrapped)
    torch.cuda.empty_cache = lambda: None


def load():
    """Returns a namespace with the reference modules on the hot path."""
    if not available():
        raise RuntimeError(f"reference not present at {REF_ROOT}")
    _install_stubs()
    _drop_cuda_device()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    import importlib

    ns = types.SimpleNamespace()
    ns.v9 = importlib.import_module("approx.approx_matmul_whole_v9")
    ns.fp8q = importlib.import_module("quantization.quantizers.fp8_quantizer")
    ns.fp_any_v5 = importlib.import_module("approx.fp_any_int_v5")
    return ns


def load_operators():
    """Reference operator classes (needs the heavier quantization package imports)."""
    load()
    import importlib

    return importlib.import_module("approx.approx_calculation")
