"""Host-side mirror of the reference's Torch-flavoured operator classes:

  QCustomLinearTorch      approx/approx_calculation.py:920-1023
  QCustomBNConv2dTorch    approx/approx_calculation.py:723-917
  QCustomConv2dTorch      approx/approx_calculation.py:551-719   (exact matmul)

Same class names, constructor keywords, method names (`approx_multiply`, `run_forward`, `im2col`,
`multiply`) and argument meaning.  What changes is underneath `approx_multiply`: the reference loops
over output columns in Python (one `custom_matmul_vectorize` call and three `.item()` syncs per
column); here ONE C-ABI call computes all columns with the biases read on the device.
The debugging prints of the reference's Linear variant (:972-984) are not reproduced.
"""
import torch
import torch.nn.functional as F
from torch import nn

from . import ops
from .quantization import BNFusedHijacker, QuantizationHijacker


class _ApproxMixin:
    """approx_multiply shared by the three operator classes (reference: :921-999, :749-814, :588-653)."""

    def _params(self, quirk):
        cap = self.custom_approx_params or {}
        key = (quirk,) + tuple(sorted((k, v) for k, v in cap.items() if not isinstance(v, (list, dict))))
        cache = self.__dict__.setdefault("_b200_param_cache", {})
        if key not in cache:
            cache[key] = ops.make_params(cap, single_column_quirk=quirk)
        return cache[key]

    def approx_multiply(self, x, y, x_bias, y_bias, res_bias):
        if y.shape[1] != 1:
            if self.approx_flag:
                return ops.approx_matmul(x, y, x_bias, y_bias, res_bias, self._params(False))
            if self.quantize_after_mult_and_add:
                q = self.res_quantizer.quantizer
                return ops.qamaa_matmul(x, y, q.n_bits, q.maxval, q.mantissa_bits, q.sign_bits)
            return x @ y
        # single output column (every depthwise group): int32-TENSOR biases in the reference
        # (:985-994) -> integer-pow numerics, no quantize_after_mult_and_add branch
        if self.approx_flag:
            return ops.approx_matmul(x, y, x_bias, y_bias, res_bias, self._params(True))
        return x @ y

    def multiply(self, x, y):
        return torch.matmul(x, y)


class QCustomLinearTorch(_ApproxMixin, QuantizationHijacker, nn.Linear):
    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:1007-1023.  Inputs with more than two dims are flattened to
        [rows, K] (the reference asserts on them)."""
        x = x.contiguous()
        weight = weight.contiguous()
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1]) if x.dim() != 2 else x
        output = self.approx_multiply(x2.detach(), weight.detach().t(), self.get_acts_fp_bias(),
                                      self.get_weights_fp_bias(), self.get_res_fp_bias())
        if bias is not None:
            output += bias
        return output.reshape(*lead, -1) if x.dim() != 2 else output


def _conv_out_hw(h, w, kh, kw, stride, padding, dilation):
    oh = (h + 2 * padding[0] - dilation[0] * (kh - 1) - 1) // stride[0] + 1
    ow = (w + 2 * padding[1] - dilation[1] * (kw - 1) - 1) // stride[1] + 1
    return oh, ow


class _ConvMixin(_ApproxMixin):
    def im2col(self, input_data, kernel_height, kernel_width, stride, padding, dilation):
        """[B,C,H,W] -> [B*Ho*Wo, C*kh*kw], K order (c,ky,kx) (approx_calculation.py:724-747)."""
        col = F.unfold(input_data, (kernel_height, kernel_width), dilation=dilation, padding=padding, stride=stride)
        return col.transpose(1, 2).reshape(-1, col.shape[1])

    def _conv_forward_cols(self, x, weight, bias, use_approx):
        x = x.contiguous().detach()
        weight = weight.contiguous().detach()
        w_b, a_b, r_b = self.get_weights_fp_bias(), self.get_acts_fp_bias(), self.get_res_fp_bias()
        B, C, H, W = x.shape
        Cout, Cg, kh, kw = weight.shape
        Ho, Wo = _conv_out_hw(H, W, kh, kw, self.stride, self.padding, self.dilation)
        col = self.im2col(x, kh, kw, self.stride, self.padding, self.dilation)
        Kg = Cg * kh * kw
        Ng = Cout // self.groups
        wcol = weight.reshape(Cout, -1)
        out = torch.empty((B * Ho * Wo, Cout), device=x.device, dtype=x.dtype)
        for g in range(self.groups):          # approx_calculation.py:858-906
            xg = col[:, g * Kg:(g + 1) * Kg]
            wg = wcol[g * Ng:(g + 1) * Ng, :]
            if use_approx:
                wb = w_b.reshape(-1)[g * Ng:(g + 1) * Ng] if w_b is not None and w_b.numel() > 1 else w_b
                out[:, g * Ng:(g + 1) * Ng] = self.approx_multiply(xg, wg.t(), a_b, wb, r_b)
            else:
                out[:, g * Ng:(g + 1) * Ng] = self.multiply(xg, wg.t())
        out = out.reshape(B, Ho, Wo, Cout).permute(0, 3, 1, 2)
        if bias is not None:
            out = out + bias.view(1, -1, 1, 1)
        return out


class QCustomBNConv2dTorch(_ConvMixin, BNFusedHijacker, nn.Conv2d):
    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:822-917."""
        return self._conv_forward_cols(x, weight, bias, use_approx=True)


class QCustomConv2dTorch(_ConvMixin, QuantizationHijacker, nn.Conv2d):
    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:655-719: conv WITHOUT BN uses the exact `multiply` (torch.matmul)."""
        return self._conv_forward_cols(x, weight, bias, use_approx=False)
