"""Host-side mirror of quantization/quantizers/fp8_quantizer.py (FP8 fake-quantizer, row A8)."""
import torch

from . import ops


def quantize_to_fp8_ste_MM(x_float, n_bits, maxval, num_mantissa_bits, sign_bits):
    """Drop-in for quantize_to_fp8_ste_MM (fp8_quantizer.py:97-173): returns (x_q, bias)."""
    if not torch.is_tensor(maxval):
        maxval = torch.tensor([float(maxval)], device=x_float.device)
    return ops.fp8_fakequant(x_float, n_bits, maxval, num_mantissa_bits, sign_bits)
