"""fp8_quantization_b200 -- B200-native (sm_100a) implementation of the approx-FP8 simulated matmul
hot path of revollllt/FP8_quantization (approx_v9 multiply + re-quantization inside
QCustomLinearTorch / QCustomBNConv2dTorch, and the FP8 fake-quantizer feeding it).

Python here is host plumbing (device memory, streams); all arithmetic runs in hand-written CUDA
kernels behind the C ABI of include/fp8approx_b200.h.  No CPU fallback, no other backend.
"""
from . import _lib, ops  # noqa: F401
from .approx_matmul import (custom_matmul_vectorize, get_error_table_NN, param_prepare,  # noqa: F401
                            float_to_fpany_absint_torch, fpany_absint_to_float_torch,
                            quant_to_fp_any_vectorize_torch)
from .fp8_quantizer import quantize_to_fp8_ste_MM  # noqa: F401

__version__ = "0.1.0"
