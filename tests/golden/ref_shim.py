"""Import shim for the UNMODIFIED reference (revollllt/FP8_quantization) on a GPU-less host.

Test infrastructure only.  Used by ``make_golden.py`` (this directory) to generate the golden
vectors committed under ``tests/golden/``.  It never travels to the GPU box: nothing in the
``-m gpu`` tests, ``smoke()`` or ``bench.py`` imports this file.

Blockers worked around:
  1. ``approx/approx_calculation.py:6``  ``import cupy``            -> empty stub module
  2. ``quantization/hijacker.py:7-8``    ``timm`` activation classes -> six empty nn.Module stubs
  3. ``approx/approx_matmul_whole_v9.py:393-549,560-589`` build tables with ``device='cuda'``
     -> ``torch.tensor`` / ``torch.zeros`` are wrapped to drop ``device='cuda'`` when CUDA is absent.
The reference sources are executed where they lie under /root/reference; nothing is copied.
"""
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("FP8_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "approx"))


def _install_stubs():
    if "cupy" not in sys.modules:
        sys.modules["cupy"] = types.ModuleType("cupy")
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
        sys.modules.update({
            "timm": timm, "timm.models": models, "timm.models.layers": layers,
            "timm.models.layers.activations": act, "timm.models.layers.activations_me": act_me,
        })


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
