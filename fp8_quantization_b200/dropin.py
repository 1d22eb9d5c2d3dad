"""Drop-in seam for the UNMODIFIED reference ("Python seam"; INTEGRATION.md section B quotes this
file verbatim).  With the reference's root on sys.path:

    import fp8_quantization_b200.dropin as b200
    b200.install()            # before or after the model is built; b200.uninstall() restores the originals

rebinds, on the reference's own classes and modules,
  approx.approx_calculation.{QCustomLinearTorch,QCustomBNConv2dTorch,QCustomConv2dTorch}.approx_multiply
      the per-column Python loop (approx_calculation.py:921-999, :749-814, :588-653) -> ONE C-ABI call
  ... .multiply                        torch.matmul (:704-705, :816-820, :1001-1003) -> b200_fq_gemm_tc (tcgen05)
  QCustomBNConv2dTorch.run_forward     im2col + group loop (:822-917) -> b200_approx_conv2d_fbias / qamaa / exact
                                       (level="run_forward" only; level="approx_multiply" keeps the reference's loop)
  quantization.quantizers.fp8_quantizer.quantize_to_fp8_ste_MM (fp8_quantizer.py:97-173) -> b200_fp8_fakequant
Signatures, return types, the custom_approx_params / run_method dicts, the hijacker forward protocol, the module
maps and every click flag stay as the reference has them.  CUDA tensors only: there is no CPU path.
"""
import importlib

from . import ops
from .approx_calculation import _ApproxMixin, _ConvMixin
from .fp8_quantizer import quantize_to_fp8_ste_MM

_saved = []


def approx_multiply(self, x, y, x_bias, y_bias, res_bias):
    """Same arguments and result as the reference method: x [Mrows,K], y [K,N] (any strides, e.g. weight.t()),
    biases = the fake-quantizers' custom_bias tensors (x_bias / res_bias may be None -> 2^(E-1)); [Mrows,N] fp32."""
    return _ApproxMixin.approx_multiply(self, x, y, x_bias, y_bias, res_bias)


def multiply(self, x, y):
    return ops.fq_matmul(x, y) if x.dim() == 2 and y.dim() == 2 else x @ y


def bn_conv_run_forward(self, x, weight, bias, offsets=None):
    return _ConvMixin._conv_forward_cols(self, x, weight, bias, use_approx=True)


def _rebind(obj, name, new):
    _saved.append((obj, name, obj.__dict__[name] if isinstance(obj, type) else getattr(obj, name)))
    setattr(obj, name, new)


def install(level="run_forward"):
    """Patches the reference modules found on sys.path (`approx`, `quantization`).  Idempotent."""
    if _saved:
        return
    assert level in ("approx_multiply", "run_forward"), level
    ac = importlib.import_module("approx.approx_calculation")
    fq = importlib.import_module("quantization.quantizers.fp8_quantizer")
    for cls in (ac.QCustomLinearTorch, ac.QCustomBNConv2dTorch, ac.QCustomConv2dTorch):
        _rebind(cls, "approx_multiply", approx_multiply)
        _rebind(cls, "multiply", multiply)
    if level == "run_forward":
        _rebind(ac.QCustomBNConv2dTorch, "run_forward", bn_conv_run_forward)
    _rebind(fq, "quantize_to_fp8_ste_MM", quantize_to_fp8_ste_MM)
    _rebind(ac, "quantize_to_fp8_ste_MM", quantize_to_fp8_ste_MM)


def uninstall():
    while _saved:
        obj, name, old = _saved.pop()
        setattr(obj, name, old)
