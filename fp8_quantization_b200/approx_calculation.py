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
from torch import nn

from . import ops
from . import quantization as _Q
from .quantization import BNFusedHijacker, QuantizationHijacker


class _ApproxMixin:
    """approx_multiply shared by the three operator classes (reference: :921-999, :749-814, :588-653).
    Written against the attributes the reference's classes carry (custom_approx_params, approx_flag,
    quantize_after_mult_and_add, res_quantizer) and called unbound, so that dropin.py can bind the very same
    functions onto the reference's own classes."""

    def _params(self, quirk):
        cap = self.custom_approx_params or {}
        key = (quirk,) + tuple(sorted((k, v) for k, v in cap.items() if not isinstance(v, (list, dict))))
        cache = self.__dict__.setdefault("_b200_param_cache", {})
        if key not in cache:
            cache[key] = ops.make_params(cap, single_column_quirk=quirk)
        return cache[key]

    def _prepared_key(self):
        """Opt-in (quantization.CACHE_FIXED_WEIGHTS): (owner, signature) under which ops keeps this layer's DECODED weight
        plane between forwards (b200_approx_params.prepared_b).  The plane is a function of the cached quantized weights
        and of the three exponent biases, i.e. of the ranges of the weight, activation and result quantizers; None
        whenever any of them is not provably fixed (then every call decodes the weights, like the reference)."""
        ent = self.__dict__.get("_b200_wcache")
        if ent is None or not _Q.CACHE_FIXED_WEIGHTS or self.training or not getattr(self, "fix_ranges_flag", False):
            return None
        sig = [ent[0]]
        for name in ("activation_quantizer", "res_quantizer"):
            q = getattr(getattr(self, name, None), "quantizer", None)
            if q is None or not hasattr(q, "_range_version") or q._bias_version != q._range_version or q.custom_bias is None:
                return None
            mv = q.maxval
            sig.append((q._range_version, mv.data_ptr(), None if mv.is_inference() else mv._version, float(q.mantissa_bits),
                        q.sign_bits))
        return (id(self), tuple(sig))

    def approx_multiply(self, x, y, x_bias, y_bias, res_bias):
        if y.shape[1] != 1:
            if self.approx_flag:
                return ops.approx_matmul(x, y, x_bias, y_bias, res_bias, _ApproxMixin._params(self, False),
                                         act_quant=self.__dict__.pop("_b200_act_quant", None),
                                         epilogue=self.__dict__.pop("_b200_ep_args", None),
                                         prepared=_ApproxMixin._prepared_key(self))
            if self.quantize_after_mult_and_add:
                q = self.res_quantizer.quantizer
                return ops.qamaa_matmul(x, y, q.n_bits, q.maxval, q.mantissa_bits, q.sign_bits, x_bias=x_bias,
                                        y_bias=y_bias)
            # `x @ y[:, i]` per column (:968-969): tcgen05.  Opt-in (quantization.enable_fusions): `+ bias` in the kernel's
            # epilogue, and the bf16 plane of cached fixed-range weights built once
            return ops.fq_matmul(x, y, bias=self.__dict__.pop("_b200_plain_bias", None),
                                 cache_b=self.__dict__.get("_b200_wcache") is not None,
                                 act_quant=self.__dict__.pop("_b200_act_quant", None),
                                 res_quant=self.__dict__.pop("_b200_res_quant", None))
        # single output column (every depthwise group): int32-TENSOR biases in the reference
        # (:985-994) -> integer-pow numerics, no quantize_after_mult_and_add branch
        if self.approx_flag:
            return ops.approx_matmul(x, y, x_bias, y_bias, res_bias, _ApproxMixin._params(self, True))
        return ops.fq_matmul(x, y)

    def multiply(self, x, y):
        """`self.multiply` = torch.matmul in the reference (:704-705); 2-D operands go to the tensor cores."""
        return ops.fq_matmul(x, y)


class QCustomLinearTorch(_ApproxMixin, QuantizationHijacker, nn.Linear):
    _plain_matmul_fuses_quantizers = True      # ops.fq_matmul(act_quant=, res_quant=): b200_fq_gemm_tc_ex

    def _fusable_input_quant(self, x):
        return self.out_features != 1          # a single output column takes the tensor-bias branch on x_q

    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:1007-1023.  Inputs with more than two dims are flattened to
        [rows, K] (the reference asserts on them)."""
        x = x.contiguous()
        weight = weight.contiguous()
        lead = x.shape[:-1]
        x2 = x.reshape(-1, x.shape[-1]) if x.dim() != 2 else x
        if bias is not None and _Q.FUSED_GEMM_EPILOGUE and not self.approx_flag and not self.quantize_after_mult_and_add \
                and weight.shape[0] != 1 and x2.is_cuda and not self.training:
            self.__dict__["_b200_plain_bias"] = bias      # plain matmul: the tensor-core kernel adds the bias itself
            bias = None
        if self.approx_flag and weight.shape[0] != 1 and "_b200_epilogue" in self.__dict__:
            # opt-in fused epilogue (quantization.FUSED_GEMM_EPILOGUE): `+ bias` and the activation inside the GEMM
            self.__dict__["_b200_ep_args"] = dict(self.__dict__.pop("_b200_epilogue"), bias=bias)
            bias = None
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
        return ops.im2col(input_data, (kernel_height, kernel_width), stride, padding, dilation)

    def _conv_forward_cols(self, x, weight, bias, use_approx):
        """run_forward of the conv operators (approx_calculation.py:822-917 / :655-719).

        approx_flag: ONE C-ABI call does im2col + group loop + approx_multiply (implicit im2col for
        groups with several output channels, the single-column kernel for depthwise groups).
        quantize_after_mult_and_add: one call as well, except single-output-channel groups, which the
        reference computes exactly (:810-811).  Otherwise: explicit im2col kernel + exact matmul."""
        x = x.contiguous().detach()
        weight = weight.contiguous().detach()
        B, C, H, W = x.shape
        Cout, Cg, kh, kw = weight.shape
        Ho, Wo = _conv_out_hw(H, W, kh, kw, self.stride, self.padding, self.dilation)
        Ng = Cout // self.groups
        if use_approx and self.approx_flag:
            ep = self.__dict__.pop("_b200_epilogue", None)        # opt-in fused epilogue: the call returns final NCHW
            out = ops.approx_conv2d(x, weight, self.get_acts_fp_bias(), self.get_weights_fp_bias(),
                                    self.get_res_fp_bias(), _ApproxMixin._params(self, Ng == 1), self.stride, self.padding,
                                    self.dilation, self.groups, act_quant=self.__dict__.pop("_b200_act_quant", None),
                                    epilogue=dict(ep, bias=bias) if ep is not None else None,
                                    prepared=_ApproxMixin._prepared_key(self))
            if ep is not None:
                return out
        elif use_approx and self.quantize_after_mult_and_add and Ng != 1:
            q = self.res_quantizer.quantizer
            out = ops.qamaa_conv2d(x, weight, q.n_bits, q.maxval, q.mantissa_bits, q.sign_bits, self.stride,
                                   self.padding, self.dilation, self.groups, x_bias=self.get_acts_fp_bias(),
                                   w_bias=self.get_weights_fp_bias())
        elif self.groups == 1:
            col = self.im2col(x, kh, kw, self.stride, self.padding, self.dilation)
            out = self.multiply(col, weight.reshape(Cout, -1).t())
        else:                                     # approx_calculation.py:858-906, exact branch, all groups at once
            out = ops.exact_grouped_conv2d(x, weight, self.stride, self.padding, self.dilation, self.groups)
        out = out.reshape(B, Ho, Wo, Cout).permute(0, 3, 1, 2)      # :909-910 (a view, like the reference)
        if bias is not None:
            out = out + bias.view(1, -1, 1, 1)
        return out


class QCustomBNConv2dTorch(_ConvMixin, BNFusedHijacker, nn.Conv2d):
    def _fusable_input_quant(self, x):
        if x.dim() != 4:
            return False
        if self.out_channels // self.groups != 1:
            return True
        # depthwise groups: the input is quantized while it is decoded into the table kernel's word plane (main mode)
        cap = self.custom_approx_params or {}
        return bool(cap.get("with_s2nn2s_opt")) and bool(cap.get("quant_btw_mult_accu", True)) and \
            not cap.get("golden_clip_OF") and not cap.get("ofuf_mode")

    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:822-917."""
        return self._conv_forward_cols(x, weight, bias, use_approx=True)


class QCustomConv2dTorch(_ConvMixin, QuantizationHijacker, nn.Conv2d):
    def run_forward(self, x, weight, bias, offsets=None):
        """approx_calculation.py:655-719: conv WITHOUT BN uses the exact `multiply` (torch.matmul)."""
        return self._conv_forward_cols(x, weight, bias, use_approx=False)
