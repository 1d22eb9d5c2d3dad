"""Tensor-level host wrappers over the C ABI (torch is used for device memory and streams only).

Every function here launches hand-written sm_100a kernels from libfp8approx_b200.so on the current
CUDA stream; none has a CPU or eager-PyTorch fallback (a CPU tensor raises).
"""
import ctypes

import torch

from . import _lib
from ._lib import ApproxParams

_WORKSPACES = {}

PARAM_KEYS = ("expo_width", "mant_width", "dnsmp_factor", "withComp", "with_approx", "with_s2nn2s_opt",
              "sim_hw_add_OFUF", "with_OF_opt", "with_UF_opt", "golden_clip_OF", "quant_btw_mult_accu",
              "debug_mode", "self_check_mode")


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _lib.B200Error("fp8_quantization_b200 has no CPU path: tensors must live on a B200 (cuda) device")


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _workspace(device, nbytes):
    key = (device.index if device.index is not None else torch.cuda.current_device())
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws


def make_params(custom_approx_params=None, *, single_column_quirk=False, allow_unpinned_format=False,
                force_generic=False, **overrides):
    """Builds the POD struct from the reference's `custom_approx_params` dict (13 keys,
    utils/click_options.py:589-606); missing keys take the defaults of custom_matmul_vectorize's
    signature (approx/approx_matmul_whole_v9.py:10-18)."""
    d = dict(expo_width=4, mant_width=3, dnsmp_factor=3, withComp=True, with_approx=True,
             with_s2nn2s_opt=False, sim_hw_add_OFUF=False, with_OF_opt=False, with_UF_opt=False,
             golden_clip_OF=False, quant_btw_mult_accu=True, debug_mode=False, self_check_mode=False)
    if custom_approx_params:
        d.update({k: v for k, v in custom_approx_params.items() if k in d})
    d.update(overrides)
    p = ApproxParams()
    p.expo_width, p.mant_width, p.dnsmp_factor = int(d["expo_width"]), int(d["mant_width"]), int(d["dnsmp_factor"])
    for k in ("withComp", "with_approx", "with_s2nn2s_opt", "sim_hw_add_OFUF", "with_OF_opt", "with_UF_opt",
              "golden_clip_OF", "quant_btw_mult_accu", "self_check_mode"):
        setattr(p, k, 1 if d[k] else 0)
    p.single_column_quirk = 1 if single_column_quirk else 0
    p.allow_unpinned_format = 1 if allow_unpinned_format else 0
    p.force_generic = 1 if force_generic else 0
    p.custom_table = None
    return p


def error_table(expo_width, mant_width, withComp, dnsmp_factor, allow_unpinned_format=False):
    """get_error_table_NN (approx/approx_matmul_whole_v9.py:555-592) as a host int8 tensor."""
    lib = _lib.load()
    n = 2 ** mant_width if 0 < mant_width <= 5 else 1
    out = torch.zeros((n, n), dtype=torch.int8)
    _lib.check(lib.b200_error_table(expo_width, mant_width, int(bool(withComp)), int(dnsmp_factor),
                                    int(allow_unpinned_format), ctypes.c_void_p(out.data_ptr())))
    return out


def approx_matmul(x, y, x_bias, y_bias, res_bias, params, custom_table=None, out=None):
    """C = approx_multiply(x [Mrows,K], y [K,N]) in ONE launch sequence.

    Biases, two forms:
      * python ints for x_bias / res_bias and an int (broadcast) or int32 tensor [N] for y_bias
        -> b200_approx_gemm;
      * the tensors the fake-quantizers produced (fp32, integer valued, any shape with 1 / N
        elements; None -> 2^(E-1) like approx_calculation.py:938-939) -> b200_approx_gemm_fbias,
        which reads them on the device: no .item() host sync.
    `y` may be any strided view (e.g. weight.t()).
    """
    lib = _lib.load()
    _require_cuda(x, y)
    assert x.dim() == 2 and y.dim() == 2
    assert x.shape[1] == y.shape[0]          # approx_matmul_whole_v9.py:20
    if x.dtype != torch.float32:
        x = x.float()
    if y.dtype != torch.float32:
        y = y.float()
    if x.stride(1) != 1:
        x = x.contiguous()
    Mrows, K = x.shape
    N = y.shape[1]
    if out is None:
        out = torch.empty((Mrows, N), dtype=torch.float32, device=x.device)
    keep = None
    if custom_table is not None:
        keep = custom_table.detach().to("cpu", torch.int8).contiguous()
        params.custom_table = keep.data_ptr()
    nbytes = lib.b200_approx_gemm_workspace_bytes(Mrows, K, N)
    ws = _workspace(x.device, nbytes)
    int_path = isinstance(x_bias, int) and isinstance(res_bias, int)
    with torch.cuda.device(x.device):
        if int_path:
            if isinstance(y_bias, int):
                y_bias = torch.full((N,), y_bias, dtype=torch.int32, device=x.device)
            else:
                y_bias = y_bias.reshape(-1).to(device=x.device, dtype=torch.int32)
                if y_bias.numel() == 1 and N > 1:
                    y_bias = y_bias.expand(N)
                assert y_bias.numel() == N
            y_bias = y_bias.contiguous()
            rc = lib.b200_approx_gemm(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                      out.stride(0), Mrows, K, N, int(x_bias), _ptr(y_bias), int(res_bias),
                                      ctypes.byref(params), _ptr(ws), ws.numel(), _stream())
        else:
            def dev_scalar(b):
                if b is None:
                    return None
                if not torch.is_tensor(b):
                    b = torch.tensor([float(b)])
                return b.reshape(-1)[:1].to(device=x.device, dtype=torch.float32)
            fa, fr = dev_scalar(x_bias), dev_scalar(res_bias)
            fb = y_bias if torch.is_tensor(y_bias) else torch.tensor([float(y_bias)])
            fb = fb.reshape(-1).to(device=x.device, dtype=torch.float32)
            if fb.numel() == 1:
                sB = 0
            else:
                if fb.numel() < N:
                    raise IndexError("index out of range: y_bias has fewer entries than output columns")
                sB = fb.stride(0)
            rc = lib.b200_approx_gemm_fbias(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                            out.stride(0), Mrows, K, N, _ptr(fa), _ptr(fb), sB, _ptr(fr),
                                            ctypes.byref(params), _ptr(ws), ws.numel(), _stream())
    if keep is not None:
        params.custom_table = None
    _lib.check(rc)
    return out


def fp8_fakequant(x_float, n_bits, maxval, num_mantissa_bits, sign_bits):
    """quantize_to_fp8_ste_MM (quantization/quantizers/fp8_quantizer.py:97-173) -> (x_q, bias)."""
    lib = _lib.load()
    _require_cuda(x_float)
    x = x_float if x_float.dtype == torch.float32 else x_float.float()
    x = x.contiguous()
    mv = maxval.to(device=x.device, dtype=torch.float32).reshape(-1).contiguous()
    mbits = float(num_mantissa_bits.item()) if torch.is_tensor(num_mantissa_bits) else float(num_mantissa_bits)
    C = mv.numel()
    if C != 1:
        # maxval.view([-1] + [1]*(x.dim()-1)): per dim-0 channel (fp8_quantizer.py:114-115)
        if x.dim() == 0 or x.shape[0] != C:
            raise RuntimeError(f"per-channel maxval of {C} entries does not match dim 0 of {tuple(x.shape)}")
        inner = x.numel() // C if C else 1
    else:
        inner = 1
    y = torch.empty_like(x)
    bias = torch.empty(C, dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        rc = lib.b200_fp8_fakequant(_ptr(x), _ptr(y), x.numel(), _ptr(mv), C, max(inner, 1), int(n_bits), mbits,
                                    int(sign_bits), _ptr(bias), _stream())
    _lib.check(rc)
    if C != 1 and maxval.dim() != x.dim():
        bias = bias.view([-1] + [1] * (x.dim() - 1))
    elif C == 1:
        bias = bias.view(maxval.shape if maxval.dim() > 0 else (1,))
    return y, bias


def fpany_encode(values, expo_width, mant_width, bias, clip_OF=False, allnorm=False, tensor_bias_quirk=False):
    lib = _lib.load()
    _require_cuda(values)
    v = values.float().contiguous()
    e = torch.empty(v.shape, dtype=torch.int32, device=v.device)
    m = torch.empty(v.shape, dtype=torch.int32, device=v.device)
    with torch.cuda.device(v.device):
        _lib.check(lib.b200_fpany_encode(_ptr(v), v.numel(), expo_width, mant_width, int(bias), int(allnorm),
                                         int(tensor_bias_quirk), int(clip_OF), _ptr(e), _ptr(m), _stream()))
    return e, m


def fpany_decode(expo, mant, expo_width, mant_width, bias, sign=None, allnorm=False):
    lib = _lib.load()
    _require_cuda(expo, mant)
    e = expo.to(torch.int32).contiguous()
    m = mant.to(torch.int32).contiguous()
    s = sign.float().contiguous() if sign is not None else None
    out = torch.empty(e.shape, dtype=torch.float32, device=e.device)
    with torch.cuda.device(e.device):
        _lib.check(lib.b200_fpany_decode(_ptr(e), _ptr(m), _ptr(s), e.numel(), expo_width, mant_width, int(bias),
                                         int(allnorm), _ptr(out), _stream()))
    return out


def fpany_quant(values, expo_width, mant_width, bias, clip_OF=True, tensor_bias_quirk=False):
    lib = _lib.load()
    _require_cuda(values)
    v = values.float().contiguous()
    out = torch.empty_like(v)
    with torch.cuda.device(v.device):
        _lib.check(lib.b200_fpany_quant(_ptr(v), v.numel(), expo_width, mant_width, int(bias), int(tensor_bias_quirk),
                                        int(clip_OF), _ptr(out), _stream()))
    return out


def kernel_launches():
    return int(_lib.load().b200_kernel_launches())


def last_path():
    return _lib.load().b200_last_path().decode()
