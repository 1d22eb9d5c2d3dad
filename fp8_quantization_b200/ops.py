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


_MBITS = {}


def _mbits(t):
    """mantissa_bits as a python float.  The reference keeps it as a one-element tensor that FPQuantizer.forward moves
    to the input's device (fp8_quantizer.py:239-240): read it once per (storage, version), not once per call."""
    if not torch.is_tensor(t):
        return float(t)
    if not t.is_cuda:
        return float(t.item())
    if t.is_inference():   # tensors made under torch.inference_mode() keep no version counter: read, do not cache
        return float(t.item())
    key = (t.data_ptr(), t._version, t.device)
    v = _MBITS.get(key)
    if v is None:
        if len(_MBITS) > 4096:
            _MBITS.clear()
        v = _MBITS[key] = float(t.item())
    return v


def _workspace(device, nbytes):
    """Scratch buffer of the calling (device, stream): it holds the device-side dispatch flags, the decoded operand
    planes and the split-K partial sums of a call, so two streams running layers concurrently must not share one."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, torch.cuda.current_stream(idx).cuda_stream)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(int(nbytes * 1.25) + 1024, dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws


_PREPARED = {}   # owner -> [signature, uint8 tensor, filled]


def clear_prepared():
    """Drops every cached decoded weight plane (see approx_matmul's `prepared`)."""
    _PREPARED.clear()


def _prepared_begin(params, prepared, nbytes, device):
    """Points `params` at the owner's prepared-B buffer (allocated / invalidated by signature); returns its record."""
    params.prepared_b, params.prepared_b_bytes, params.prepared_b_ready = None, 0, 0
    if prepared is None or nbytes == 0:
        return None
    owner, sig = prepared
    ent = _PREPARED.get(owner)
    if ent is None or ent[0] != sig or ent[1].numel() < nbytes or ent[1].device != device:
        ent = [sig, torch.empty(int(nbytes), dtype=torch.uint8, device=device), False]
        _PREPARED[owner] = ent
    params.prepared_b, params.prepared_b_bytes = ent[1].data_ptr(), ent[1].numel()
    params.prepared_b_ready = 1 if ent[2] else 0
    return ent


def _prepared_end(params, ent, ok):
    params.prepared_b, params.prepared_b_bytes, params.prepared_b_ready = None, 0, 0
    # (a fill that was only CAPTURED into a CUDA graph has not run yet: the buffer counts as filled only after an eager call)
    if ent is not None and ok and not torch.cuda.is_current_stream_capturing():
        ent[2] = True


def table_variant(device=None):
    """Which variant of the table GEMM computed the last approx / qamaa call on this (device, stream): "byte-cells"
    (int8 cells in shared memory, integer accumulators; picked on the device for mant_width >= 4 when every sum is
    exact and escapes are rare) or "16-bit".  Reads one flag word of the workspace: synchronises (tests, tools)."""
    idx = torch.cuda.current_device() if device is None else torch.device(device).index
    ws = _WORKSPACES.get((idx, torch.cuda.current_stream(idx).cuda_stream))
    if ws is None:
        return None
    flags = ws[:64].view(torch.int32).cpu()
    return {"variant": "byte-cells" if int(flags[11]) else "16-bit", "escapes": int(flags[8]) & 0xFFFFFFFF,
            "rare_b": int(flags[13]) & 0xFFFFFFFF, "tmax": (int(flags[4]), int(flags[5])), "tb_eff": int(flags[12]),
            "offgrid": int(flags[0])}


def make_params(custom_approx_params=None, *, single_column_quirk=False, allow_unpinned_format=False,
                force_generic=False, ofuf_mode=None, force_table16=False, **overrides):
    """Builds the POD struct from the reference's `custom_approx_params` dict (13 keys,
    utils/click_options.py:589-606); missing keys take the defaults of custom_matmul_vectorize's
    signature (approx/approx_matmul_whole_v9.py:10-18).
    ofuf_mode: 0 = approx_v9 (sim_hw_add_OFUF / with_OF_opt / with_UF_opt ignored, the reference's behaviour);
    1 or "v5" = the v5 integer-add GEMM in which those flags are live.  Opt-in only: taken
    from the argument, else from an extra `ofuf_mode` key of the dict (not one of the reference's 13 keys)."""
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
    p.force_table16 = 1 if force_table16 else 0    # tests / before-after runs: skip the byte-cell table variant
    if ofuf_mode is None:
        ofuf_mode = (custom_approx_params or {}).get("ofuf_mode", 0)
    p.ofuf_mode = 1 if ofuf_mode in (1, True, "v5") else 0
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


def comp_table_v5(expo_width, mant_width, withComp, dnsmp_factor):
    """get_comp_table_NN of the v5 integer-add multiplier (approx/approx_matmul_whole_v5.py:516-542), host int8."""
    lib = _lib.load()
    n = 2 ** mant_width if 0 < mant_width <= 5 else 1
    out = torch.zeros((n, n), dtype=torch.int8)
    _lib.check(lib.b200_comp_table_v5(expo_width, mant_width, int(bool(withComp)), int(dnsmp_factor),
                                      ctypes.c_void_p(out.data_ptr())))
    return out


def _act_quant_struct(act_quant, device):
    """(maxval [1], bias [1], mantissa_bits, n_bits, sign_bits) -> (struct b200_act_quant, tensors kept alive)."""
    maxval, bias, mbits, n_bits, sign_bits = act_quant
    mv = maxval.to(device=device, dtype=torch.float32).reshape(-1)
    bs = bias.to(device=device, dtype=torch.float32).reshape(-1)
    if mv.numel() != 1 or bs.numel() != 1:
        raise ValueError("the fused input quantizer needs a per-tensor range")
    aq = _lib.ActQuant(mv.data_ptr(), bs.data_ptr(), float(mbits), int(n_bits), int(sign_bits))
    return aq, (mv, bs)


def _epilogue_struct(epilogue, n_channels, device):
    """dict(bias=, bn=(mean, var, gamma, beta, eps), act=0|1|2) -> (struct b200_epilogue, tensors kept alive).
    Every per-channel tensor must have n_channels elements."""
    keep = []

    def vec(t):
        if t is None:
            return None
        t = t.detach().to(device=device, dtype=torch.float32).reshape(-1).contiguous()
        if t.numel() != n_channels:
            raise ValueError(f"epilogue: expected {n_channels} per-channel values, got {t.numel()}")
        keep.append(t)
        return t.data_ptr()
    ep = _lib.Epilogue()
    ep.bias = vec(epilogue.get("bias"))
    bn = epilogue.get("bn")
    if bn is not None:
        mean, var, gamma, beta, eps = bn
        ep.bn_mean, ep.bn_var, ep.bn_gamma, ep.bn_beta = vec(mean), vec(var), vec(gamma), vec(beta)
        ep.bn_eps = float(eps)
    ep.act = int(epilogue.get("act", 0))
    return ep, keep


def approx_matmul(x, y, x_bias, y_bias, res_bias, params, custom_table=None, out=None, act_quant=None, epilogue=None,
                  prepared=None):
    """C = approx_multiply(x [Mrows,K], y [K,N]) in ONE launch sequence.

    Biases, two forms:
      * python ints for x_bias / res_bias and an int (broadcast) or int32 tensor [N] for y_bias
        -> b200_approx_gemm;
      * the tensors the fake-quantizers produced (fp32, integer valued, any shape with 1 / N
        elements; None -> 2^(E-1) like approx_calculation.py:938-939) -> b200_approx_gemm_fbias,
        which reads them on the device: no .item() host sync.
    `y` may be any strided view (e.g. weight.t()).
    `act_quant` = (maxval, bias, mantissa_bits, n_bits, sign_bits): `x` holds RAW activations and the input
    fake-quantizer runs inside the operand decode pass (b200_approx_gemm_fbias_fq); bit-identical to quantizing first.
    `epilogue` = dict(bias=[N], bn=(mean, var, gamma, beta, eps), act=0|1|2), all optional (opt-in):
    `+ bias`, eval batch norm and ReLU / ReLU6 evaluated in the GEMM's epilogue (b200_approx_gemm_fbias_ep).
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
    if (act_quant is not None or epilogue is not None) and int_path:
        raise ValueError("the fused input quantizer / epilogue need the device-bias form of the call")
    # `prepared` = (owner, signature), opt-in: the decoded weight plane is kept between calls with the same signature
    pent = _prepared_begin(params, prepared, lib.b200_approx_gemm_prepared_b_bytes(K, N) if prepared is not None else 0,
                           x.device)
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
            if epilogue is not None:
                aq, alive = _act_quant_struct(act_quant, x.device) if act_quant is not None else (None, None)
                ep, alive2 = _epilogue_struct(epilogue, N, x.device)
                rc = lib.b200_approx_gemm_fbias_ep(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                                   out.stride(0), Mrows, K, N, _ptr(fa), _ptr(fb), sB, _ptr(fr),
                                                   ctypes.byref(params), ctypes.byref(aq) if aq is not None else None,
                                                   ctypes.byref(ep), _ptr(ws), ws.numel(), _stream())
            elif act_quant is not None:
                aq, alive = _act_quant_struct(act_quant, x.device)
                rc = lib.b200_approx_gemm_fbias_fq(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                                   out.stride(0), Mrows, K, N, _ptr(fa), _ptr(fb), sB, _ptr(fr),
                                                   ctypes.byref(params), ctypes.byref(aq), _ptr(ws), ws.numel(), _stream())
            else:
                rc = lib.b200_approx_gemm_fbias(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                                out.stride(0), Mrows, K, N, _ptr(fa), _ptr(fb), sB, _ptr(fr),
                                                ctypes.byref(params), _ptr(ws), ws.numel(), _stream())
    if keep is not None:
        params.custom_table = None
    _prepared_end(params, pent, rc == 0)
    _lib.check(rc)
    return out


def _dev_scalar(b, device):
    if b is None:
        return None
    if not torch.is_tensor(b):
        b = torch.tensor([float(b)])
    return b.reshape(-1)[:1].to(device=device, dtype=torch.float32)


def _bias_vec(b, n, device, what):
    """Per-column exponent-bias hint as (device fp32 vector, stride) -- stride 0 broadcasts one value."""
    fb = b if torch.is_tensor(b) else torch.tensor([float(b)])
    fb = fb.reshape(-1).to(device=device, dtype=torch.float32)
    if fb.numel() == 1:
        return fb, 0
    if fb.numel() < n:
        raise IndexError(f"index out of range: {what} has fewer entries than output columns")
    return fb, fb.stride(0)


def qamaa_matmul(x, y, n_bits, maxval, mantissa_bits, sign_bits, out=None, x_bias=None, y_bias=None):
    """quantize_after_mult_and_add column loop (approx/approx_calculation.py:959-967) in ONE launch:
    C = Q(sum_k Q(x[:,k] * y[k,:])) with Q = quantize_to_fp8_ste_MM(., n_bits, maxval, mantissa_bits,
    sign_bits)[0]; `maxval` is the result quantizer's per-tensor range (read on the device)."""
    lib = _lib.load()
    _require_cuda(x, y)
    assert x.dim() == 2 and y.dim() == 2 and x.shape[1] == y.shape[0]
    x = x.float() if x.dtype != torch.float32 else x
    y = y.float() if y.dtype != torch.float32 else y
    if x.stride(1) != 1:
        x = x.contiguous()
    mv = torch.as_tensor(maxval).to(device=x.device, dtype=torch.float32).reshape(-1)
    if mv.numel() != 1:
        raise ValueError("quantize_after_mult_and_add needs a per-tensor result range")
    mbits = _mbits(mantissa_bits)
    Mrows, K = x.shape
    N = y.shape[1]
    if out is None:
        out = torch.empty((Mrows, N), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        if y_bias is not None:
            # operand exponent biases known (the fake-quantizers' custom_bias): table kernel
            fb, sB = _bias_vec(y_bias, N, x.device, "y_bias")
            fa = _dev_scalar(x_bias, x.device)
            ws = _workspace(x.device, lib.b200_qamaa_gemm_workspace_bytes(Mrows, K, N))
            rc = lib.b200_qamaa_gemm_fbias(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out),
                                           out.stride(0), Mrows, K, N, _ptr(mv), int(n_bits), mbits, int(sign_bits),
                                           _ptr(fa), _ptr(fb), sB, _ptr(ws), ws.numel(), _stream())
        else:
            rc = lib.b200_qamaa_gemm(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out), out.stride(0),
                                     Mrows, K, N, _ptr(mv), int(n_bits), mbits, int(sign_bits), _stream())
    _lib.check(rc)
    return out


_BPLANES = {}


def _prepared_b(lib, y, K, N):
    """bf16 plane of a right operand that stays the same call after call (b200_fq_gemm_tc_prepare_b), built once per
    (storage, version, view).  The entry holds `y` itself, so its storage cannot be freed and reused under the key."""
    key = (y.data_ptr(), y._version if not y.is_inference() else -1, tuple(y.shape), tuple(y.stride()), y.device)
    ent = _BPLANES.get(key)
    if ent is None:
        if len(_BPLANES) >= 512:
            _BPLANES.clear()
        nbytes = lib.b200_fq_gemm_tc_bplane_bytes(K, N)
        plane = torch.empty(nbytes, dtype=torch.uint8, device=y.device)
        _lib.check(lib.b200_fq_gemm_tc_prepare_b(_ptr(y), y.stride(0), y.stride(1), K, N, _ptr(plane), nbytes, _stream()))
        ent = _BPLANES[key] = (plane, y)
    return ent[0]


def fq_matmul(x, y, bias=None, out=None, cache_b=False, act_quant=None, res_quant=None):
    """The non-approx `x @ y` of fake-quantized operands (approx/approx_calculation.py:968-969, :704-705) on the
    tcgen05 tensor cores: bf16 operands (exact for FP8-grid values), fp32 accumulation in tensor memory.
    Operands that are not exact in bf16 are detected on the device and computed in fp32 instead.
    bias: [N], added to every row in the kernel's epilogue (`output += bias`, :1020-1021).
    cache_b (opt-in): `y` is a tensor that stays the same call after call (fixed-range weights): its bf16 plane is
    built once and reused (b200_fq_gemm_tc_pre).
    act_quant / res_quant (opt-in) = (maxval, bias, mantissa_bits, n_bits, sign_bits): the input fake-quantizer runs while
    `x` (then RAW activations) is converted to bf16, the result fake-quantizer in the epilogue after the bias
    (b200_fq_gemm_tc_ex); bit-identical to the separate passes."""
    lib = _lib.load()
    _require_cuda(x, y)
    assert x.dim() == 2 and y.dim() == 2 and x.shape[1] == y.shape[0]
    x = x.float() if x.dtype != torch.float32 else x
    y = y.float() if y.dtype != torch.float32 else y
    if x.stride(1) != 1:
        x = x.contiguous()
    Mrows, K = x.shape
    N = y.shape[1]
    if out is None:
        out = torch.empty((Mrows, N), dtype=torch.float32, device=x.device)
    b = None
    if bias is not None:
        b = bias.detach().to(device=x.device, dtype=torch.float32).reshape(-1).contiguous()
        assert b.numel() == N
    ws = _workspace(x.device, lib.b200_fq_gemm_tc_workspace_bytes(Mrows, K, N))
    with torch.cuda.device(x.device):
        if act_quant is not None or res_quant is not None:
            plane = _prepared_b(lib, y, K, N) if (cache_b and K > 0 and N > 0 and Mrows > 0) else None
            aqi, keep1 = _act_quant_struct(act_quant, x.device) if act_quant is not None else (None, None)
            aqo, keep2 = _act_quant_struct(res_quant, x.device) if res_quant is not None else (None, None)
            rc = lib.b200_fq_gemm_tc_ex(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(plane),
                                        ctypes.byref(aqi) if aqi is not None else None,
                                        ctypes.byref(aqo) if aqo is not None else None, _ptr(out), out.stride(0), Mrows, K, N,
                                        _ptr(b), _ptr(ws), ws.numel(), _stream())
        elif cache_b and K > 0 and N > 0 and Mrows > 0:
            plane = _prepared_b(lib, y, K, N)
            rc = lib.b200_fq_gemm_tc_pre(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(plane), _ptr(out),
                                         out.stride(0), Mrows, K, N, _ptr(b), _ptr(ws), ws.numel(), _stream())
        else:
            rc = lib.b200_fq_gemm_tc(_ptr(x), x.stride(0), _ptr(y), y.stride(0), y.stride(1), _ptr(out), out.stride(0),
                                     Mrows, K, N, _ptr(b), _ptr(ws), ws.numel(), _stream())
    _lib.check(rc)
    return out


def conv_geom(x_shape, w_shape, stride, padding, dilation, groups):
    g = _lib.ConvGeom()
    g.batch, g.C_in, g.H, g.W = (int(v) for v in x_shape)
    g.C_out, _, g.kh, g.kw = (int(v) for v in w_shape)
    g.stride_h, g.stride_w = (int(v) for v in stride)
    g.pad_h, g.pad_w = (int(v) for v in padding)
    g.dil_h, g.dil_w = (int(v) for v in dilation)
    g.groups = int(groups)
    return g


def conv_out_shape(g):
    lib = _lib.load()
    rows = ctypes.c_int64()
    ho, wo, kg, ng = (ctypes.c_int32() for _ in range(4))
    _lib.check(lib.b200_conv_out_shape(ctypes.byref(g), ctypes.byref(rows), ctypes.byref(ho), ctypes.byref(wo),
                                       ctypes.byref(kg), ctypes.byref(ng)))
    return rows.value, ho.value, wo.value, kg.value, ng.value


def im2col(x, kernel_size, stride, padding, dilation):
    """[B,C,H,W] -> [B*Ho*Wo, C*kh*kw], K order (c,ky,kx) (approx/approx_calculation.py:724-747)."""
    lib = _lib.load()
    _require_cuda(x)
    x = x.float().contiguous()
    g = conv_geom(x.shape, (1, x.shape[1], kernel_size[0], kernel_size[1]), stride, padding, dilation, 1)
    rows, _, _, kg, _ = conv_out_shape(g)
    out = torch.empty((rows, kg), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.b200_im2col(_ptr(x), _ptr(out), ctypes.byref(g), _stream()))
    return out


def exact_grouped_conv2d(x, weight, stride, padding, dilation, groups):
    """Exact branch of the conv group loop (approx/approx_calculation.py:796-797, :810-811) for groups > 1:
    every group in one launch over an implicit im2col view.  Returns [B*Ho*Wo, C_out] fp32."""
    lib = _lib.load()
    _require_cuda(x, weight)
    x = x.float().contiguous()
    weight = weight.float().contiguous()
    g = conv_geom(x.shape, weight.shape, stride, padding, dilation, groups)
    rows, _, _, _, _ = conv_out_shape(g)
    out = torch.empty((rows, g.C_out), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.b200_exact_grouped_conv2d(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _stream()))
    return out


def approx_conv2d(x, weight, x_bias, w_bias, res_bias, params, stride, padding, dilation, groups, act_quant=None,
                  epilogue=None, prepared=None):
    """im2col + group loop + approx_multiply of QCustomBNConv2dTorch.run_forward
    (approx/approx_calculation.py:846-906) in one C-ABI call.  Returns [B*Ho*Wo, C_out] fp32.
    Biases are the fake-quantizers' fp32 tensors (or None -> 2^(E-1)); read on the device.
    `epilogue` (opt-in, see approx_matmul): the call then returns the FINAL tensor, contiguous NCHW
    [B, C_out, Ho, Wo] = act(bn(conv + bias)) (b200_approx_conv2d_fbias_ep)."""
    lib = _lib.load()
    _require_cuda(x, weight)
    x = x.float().contiguous()
    weight = weight.float().contiguous()
    g = conv_geom(x.shape, weight.shape, stride, padding, dilation, groups)
    rows, ho, wo, _, _ = conv_out_shape(g)
    if epilogue is not None:
        out = torch.empty((g.batch, g.C_out, ho, wo), dtype=torch.float32, device=x.device)
    else:
        out = torch.empty((rows, g.C_out), dtype=torch.float32, device=x.device)
    fa, fr = _dev_scalar(x_bias, x.device), _dev_scalar(res_bias, x.device)
    fb = w_bias if torch.is_tensor(w_bias) else torch.tensor([float(w_bias if w_bias is not None else
                                                                    2 ** (params.expo_width - 1))])
    fb = fb.reshape(-1).to(device=x.device, dtype=torch.float32)
    if fb.numel() == 1:
        sB = 0
    else:
        if fb.numel() != g.C_out:
            raise IndexError("weight fp bias does not have one entry per output channel")
        sB = fb.stride(0)
    nbytes = lib.b200_approx_conv2d_workspace_bytes(ctypes.byref(g))
    ws = _workspace(x.device, nbytes) if nbytes else None
    pent = _prepared_begin(params, prepared,
                           lib.b200_approx_conv2d_prepared_b_bytes(ctypes.byref(g)) if prepared is not None else 0, x.device)
    with torch.cuda.device(x.device):
        if epilogue is not None:
            aq, alive = _act_quant_struct(act_quant, x.device) if act_quant is not None else (None, None)
            ep, alive2 = _epilogue_struct(epilogue, g.C_out, x.device)
            rc = lib.b200_approx_conv2d_fbias_ep(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _ptr(fa), _ptr(fb), sB,
                                                 _ptr(fr), ctypes.byref(params),
                                                 ctypes.byref(aq) if aq is not None else None, ctypes.byref(ep), _ptr(ws),
                                                 ws.numel() if ws is not None else 0, _stream())
        elif act_quant is not None:
            aq, alive = _act_quant_struct(act_quant, x.device)
            rc = lib.b200_approx_conv2d_fbias_fq(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _ptr(fa), _ptr(fb), sB,
                                                 _ptr(fr), ctypes.byref(params), ctypes.byref(aq), _ptr(ws),
                                                 ws.numel() if ws is not None else 0, _stream())
        else:
            rc = lib.b200_approx_conv2d_fbias(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _ptr(fa), _ptr(fb), sB,
                                              _ptr(fr), ctypes.byref(params), _ptr(ws), ws.numel() if ws is not None else 0,
                                              _stream())
    _prepared_end(params, pent, rc == 0)
    _lib.check(rc)
    return out


def qamaa_conv2d(x, weight, n_bits, maxval, mantissa_bits, sign_bits, stride, padding, dilation, groups,
                 x_bias=None, w_bias=None):
    """quantize_after_mult_and_add conv (approx/approx_calculation.py:787-795 in the group loop)."""
    lib = _lib.load()
    _require_cuda(x, weight)
    x = x.float().contiguous()
    weight = weight.float().contiguous()
    g = conv_geom(x.shape, weight.shape, stride, padding, dilation, groups)
    rows, _, _, _, _ = conv_out_shape(g)
    out = torch.empty((rows, g.C_out), dtype=torch.float32, device=x.device)
    mv = torch.as_tensor(maxval).to(device=x.device, dtype=torch.float32).reshape(-1)
    if mv.numel() != 1:
        raise ValueError("quantize_after_mult_and_add needs a per-tensor result range")
    mbits = _mbits(mantissa_bits)
    with torch.cuda.device(x.device):
        if w_bias is not None:
            fb, sB = _bias_vec(w_bias, g.C_out, x.device, "weight fp bias")
            fa = _dev_scalar(x_bias, x.device)
            ws = _workspace(x.device, lib.b200_qamaa_conv2d_workspace_bytes(ctypes.byref(g)))
            rc = lib.b200_qamaa_conv2d_fbias(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _ptr(mv), int(n_bits),
                                             mbits, int(sign_bits), _ptr(fa), _ptr(fb), sB, _ptr(ws), ws.numel(),
                                             _stream())
        else:
            rc = lib.b200_qamaa_conv2d(_ptr(x), _ptr(weight), _ptr(out), ctypes.byref(g), _ptr(mv), int(n_bits), mbits,
                                       int(sign_bits), _stream())
    _lib.check(rc)
    return out


def fp8_fakequant(x_float, n_bits, maxval, num_mantissa_bits, sign_bits):
    """quantize_to_fp8_ste_MM (quantization/quantizers/fp8_quantizer.py:97-173) -> (x_q, bias)."""
    lib = _lib.load()
    _require_cuda(x_float)
    x = x_float if x_float.dtype == torch.float32 else x_float.float()
    x = x.contiguous()
    mv = maxval.to(device=x.device, dtype=torch.float32).reshape(-1).contiguous()
    mbits = _mbits(num_mantissa_bits)
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


def approx_mult_v5(x_int, y_int, B_neg, x_mant, y_mant, comp_table_NN, sim_hw_add_OFUF, with_OF_opt, with_UF_opt,
                   expo_width, mant_width):
    """approx_mult_new (approx/approx_matmul_whole_v5.py:155-182): the integer-add approximate multiply with
    the simulated hardware-adder overflow / underflow -- the only live definition of sim_hw_add_OFUF,
    with_OF_opt and with_UF_opt (approx_v9 ignores them).  int32 code tensors in, int32 codes out."""
    lib = _lib.load()
    _require_cuda(x_int, y_int)
    shape = torch.broadcast_shapes(x_int.shape, y_int.shape)
    xi = x_int.to(torch.int32).expand(shape).contiguous()
    yi = y_int.to(torch.int32).expand(shape).contiguous()
    xm = x_mant.to(torch.int32).expand(shape).contiguous() if x_mant is not None else None
    ym = y_mant.to(torch.int32).expand(shape).contiguous() if y_mant is not None else None
    t = comp_table_NN.to(device=xi.device, dtype=torch.int8).contiguous() if comp_table_NN is not None else None
    out = torch.empty(shape, dtype=torch.int32, device=xi.device)
    with torch.cuda.device(xi.device):
        _lib.check(lib.b200_approx_mult_v5(_ptr(xi), _ptr(yi), _ptr(xm), _ptr(ym), _ptr(t), xi.numel(), int(B_neg),
                                           int(expo_width), int(mant_width), int(bool(sim_hw_add_OFUF)),
                                           int(bool(with_OF_opt)), int(bool(with_UF_opt)), _ptr(out), _stream()))
    return out


def bn_act(x, running_mean, running_var, gamma, beta, eps, act=0):
    """Eval-mode BatchNorm + activation (quantization/quantized_folded_bn.py:66-78) with ATen's CPU op sequence
    (bit-identical to the reference on CPU tensors).  x: [B,C,H,W] in any layout whose (h,w) plane is linear
    (contiguous or channels-last); returns a contiguous NCHW tensor.  act: 0 none, 1 ReLU, 2 ReLU6."""
    lib = _lib.load()
    _require_cuda(x)
    assert x.dim() == 4
    if x.dtype != torch.float32:
        x = x.float()
    B, C, H, W = x.shape
    if W > 1 and H > 1 and x.stride(2) != W * x.stride(3):
        x = x.contiguous()
    sp = x.stride(3) if W > 1 else (x.stride(2) if H > 1 else 1)
    y = torch.empty((B, C, H, W), dtype=torch.float32, device=x.device)
    f = lambda t: None if t is None else t.detach().to(device=x.device, dtype=torch.float32).contiguous()
    m, v, g, b = f(running_mean), f(running_var), f(gamma), f(beta)
    with torch.cuda.device(x.device):
        _lib.check(lib.b200_bn_act(_ptr(x), x.stride(0), x.stride(1), sp, _ptr(y), B, C, H * W, _ptr(m), _ptr(v),
                                   _ptr(g), _ptr(b), float(eps), int(act), _stream()))
    return y


def set_autotune(on):
    """Measured (tile, split-K) choice of the table GEMM (b200_set_autotune).  OFF by default in the C ABI -- no
    entry point ever synchronises with the host -- and switched on explicitly by the callers that want it (bench.py,
    the eval harness, tools/): the first call per problem shape then times a few configurations (host syncs, never
    inside CUDA-graph capture).  Every configuration computes identical bits.  Returns the previous setting."""
    return bool(_lib.load().b200_set_autotune(1 if on else 0))


def kernel_launches():
    return int(_lib.load().b200_kernel_launches())


def last_path():
    return _lib.load().b200_last_path().decode()
