"""ctypes binding of libfp8approx_b200.so (include/fp8approx_b200.h).

There is NO CPU fallback and no alternative backend: if the shared library is missing, or no sm_100
device is visible when a compute entry point is called, the call raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfp8approx_b200.so")

B200_OK = 0
ERR_INVALID_FORMAT = -1
ERR_UNBOUND_TABLE = -2
ERR_SHAPE = -3
ERR_ARGUMENT = -4
ERR_WORKSPACE = -5
ERR_CUDA = -6
ERR_NO_DEVICE = -7

c_int = ctypes.c_int
c_i64 = ctypes.c_int64
c_u8 = ctypes.c_uint8
c_f32 = ctypes.c_float
c_vp = ctypes.c_void_p
c_sz = ctypes.c_size_t


class ApproxParams(ctypes.Structure):
    """struct b200_approx_params (custom_approx_params, utils/click_options.py:589-606)."""
    _fields_ = [
        ("expo_width", ctypes.c_int32), ("mant_width", ctypes.c_int32), ("dnsmp_factor", ctypes.c_int32),
        ("withComp", c_u8), ("with_approx", c_u8), ("with_s2nn2s_opt", c_u8), ("sim_hw_add_OFUF", c_u8),
        ("with_OF_opt", c_u8), ("with_UF_opt", c_u8), ("golden_clip_OF", c_u8), ("quant_btw_mult_accu", c_u8),
        ("self_check_mode", c_u8), ("single_column_quirk", c_u8), ("allow_unpinned_format", c_u8),
        ("force_generic", c_u8), ("ofuf_mode", c_u8), ("force_table16", c_u8), ("prepared_b_ready", c_u8),
        ("custom_table", c_vp), ("prepared_b", c_vp), ("prepared_b_bytes", ctypes.c_uint64),
    ]


class ActQuant(ctypes.Structure):
    """struct b200_act_quant (include/fp8approx_b200.h): the fused input fake-quantizer."""
    _fields_ = [("maxval_dev", ctypes.c_void_p), ("bias_dev", ctypes.c_void_p), ("mantissa_bits", ctypes.c_float),
                ("n_bits", ctypes.c_int32), ("sign_bits", ctypes.c_int32)]


class Epilogue(ctypes.Structure):
    """struct b200_epilogue (include/fp8approx_b200.h): fused +bias -> eval BN -> activation -> store."""
    _fields_ = [("bias", ctypes.c_void_p), ("bn_mean", ctypes.c_void_p), ("bn_var", ctypes.c_void_p),
                ("bn_gamma", ctypes.c_void_p), ("bn_beta", ctypes.c_void_p), ("bn_eps", ctypes.c_float),
                ("act", ctypes.c_int32)]


class ConvGeom(ctypes.Structure):
    """struct b200_conv_geom (conv geometry of QCustomBNConv2dTorch.run_forward, approx_calculation.py:822-857)."""
    _fields_ = [("batch", ctypes.c_int64), ("C_in", ctypes.c_int32), ("H", ctypes.c_int32), ("W", ctypes.c_int32),
                ("C_out", ctypes.c_int32), ("kh", ctypes.c_int32), ("kw", ctypes.c_int32),
                ("stride_h", ctypes.c_int32), ("stride_w", ctypes.c_int32), ("pad_h", ctypes.c_int32),
                ("pad_w", ctypes.c_int32), ("dil_h", ctypes.c_int32), ("dil_w", ctypes.c_int32),
                ("groups", ctypes.c_int32)]


# every symbol include/fp8approx_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    "b200_version": (ctypes.c_char_p, []),
    "b200_last_error": (ctypes.c_char_p, []),
    "b200_kernel_launches": (ctypes.c_uint64, []),
    "b200_last_path": (ctypes.c_char_p, []),
    "b200_error_table": (c_int, [c_int, c_int, c_int, c_int, c_int, c_vp]),
    "b200_comp_table_v5": (c_int, [c_int, c_int, c_int, c_int, c_vp]),
    "b200_fp8_fakequant": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_int, c_f32, c_int, c_vp, c_vp]),
    "b200_fpany_encode": (c_int, [c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp, c_vp]),
    "b200_fpany_decode": (c_int, [c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "b200_fpany_quant": (c_int, [c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_vp, c_vp]),
    "b200_approx_mult_v5": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_int, c_int, c_int, c_int, c_int, c_int, c_vp,
                                    c_vp]),
    "b200_approx_gemm_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64]),
    "b200_approx_gemm": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_vp,
                                 c_int, ctypes.POINTER(ApproxParams), c_vp, c_sz, c_vp]),
    "b200_approx_gemm_fbias": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp,
                                       c_i64, c_vp, ctypes.POINTER(ApproxParams), c_vp, c_sz, c_vp]),
    "b200_approx_gemm_fbias_fq": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp,
                                          c_i64, c_vp, ctypes.POINTER(ApproxParams), ctypes.POINTER(ActQuant), c_vp, c_sz,
                                          c_vp]),
    "b200_approx_conv2d_fbias_fq": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp, c_vp, c_i64, c_vp,
                                            ctypes.POINTER(ApproxParams), ctypes.POINTER(ActQuant), c_vp, c_sz, c_vp]),
    "b200_approx_gemm_fbias_ep": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp,
                                          c_i64, c_vp, ctypes.POINTER(ApproxParams), ctypes.POINTER(ActQuant),
                                          ctypes.POINTER(Epilogue), c_vp, c_sz, c_vp]),
    "b200_approx_conv2d_fbias_ep": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp, c_vp, c_i64, c_vp,
                                            ctypes.POINTER(ApproxParams), ctypes.POINTER(ActQuant),
                                            ctypes.POINTER(Epilogue), c_vp, c_sz, c_vp]),
    "b200_qamaa_gemm": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_int, c_f32,
                                c_int, c_vp]),
    "b200_qamaa_gemm_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64]),
    "b200_qamaa_gemm_fbias": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_int,
                                      c_f32, c_int, c_vp, c_vp, c_i64, c_vp, c_sz, c_vp]),
    "b200_static_scratch_bytes": (c_sz, []),
    "b200_approx_gemm_prepared_b_bytes": (c_sz, [c_i64, c_i64]),
    "b200_qamaa_gemm_fbias_static": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp,
                                             c_int, c_f32, c_int, c_vp, c_vp, c_i64, c_vp]),
    "b200_qamaa_conv2d_workspace_bytes": (c_sz, [ctypes.POINTER(ConvGeom)]),
    "b200_qamaa_conv2d_fbias": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp, c_int, c_f32, c_int, c_vp,
                                        c_vp, c_i64, c_vp, c_sz, c_vp]),
    "b200_fq_gemm_tc_workspace_bytes": (c_sz, [c_i64, c_i64, c_i64]),
    "b200_fq_gemm_tc": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_sz,
                                c_vp]),
    "b200_fq_gemm_tc_bplane_bytes": (c_sz, [c_i64, c_i64]),
    "b200_fq_gemm_tc_prepare_b": (c_int, [c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_sz, c_vp]),
    "b200_fq_gemm_tc_pre": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_sz,
                                    c_vp]),
    "b200_fq_gemm_tc_ex": (c_int, [c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, ctypes.POINTER(ActQuant), ctypes.POINTER(ActQuant),
                                   c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_sz, c_vp]),
    "b200_conv_out_shape": (c_int, [ctypes.POINTER(ConvGeom), ctypes.POINTER(c_i64), ctypes.POINTER(ctypes.c_int32),
                                    ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32),
                                    ctypes.POINTER(ctypes.c_int32)]),
    "b200_im2col": (c_int, [c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp]),
    "b200_exact_grouped_conv2d": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp]),
    "b200_approx_conv2d_workspace_bytes": (c_sz, [ctypes.POINTER(ConvGeom)]),
    "b200_approx_conv2d_prepared_b_bytes": (c_sz, [ctypes.POINTER(ConvGeom)]),
    "b200_approx_conv2d_fbias": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp, c_vp, c_i64, c_vp,
                                         ctypes.POINTER(ApproxParams), c_vp, c_sz, c_vp]),
    "b200_qamaa_conv2d": (c_int, [c_vp, c_vp, c_vp, ctypes.POINTER(ConvGeom), c_vp, c_int, c_f32, c_int, c_vp]),
    "b200_bn_act": (c_int, [c_vp, c_i64, c_i64, c_i64, c_vp, c_i64, c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_f32, c_int,
                            c_vp]),
    "b200_set_autotune": (c_int, [c_int]),
    "b200_autotune_runs": (ctypes.c_uint64, []),
    "b200_int_peak_probe": (c_int, [c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
    "b200_lds_peak_probe": (c_int, [c_int, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
    "b200_profile_begin": (c_int, []),
    "b200_profile_end": (c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(c_int)]),
}

_lib = None


class B200Error(RuntimeError):
    pass


def load():
    """Loads the shared library (once).  Raises if it has not been built: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C fp8_quantization_b200/csrc`. There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    return load().b200_last_error().decode()


def check(rc):
    """Maps a return code to the exception the reference raises at the same condition."""
    if rc == B200_OK:
        return
    msg = last_error()
    if rc == ERR_INVALID_FORMAT:
        raise ValueError(msg)                      # approx_matmul_whole_v9.py:590
    if rc == ERR_UNBOUND_TABLE:
        raise UnboundLocalError(msg)               # approx_matmul_whole_v9.py:566-586
    if rc == ERR_SHAPE:
        raise AssertionError(msg)                  # approx_matmul_whole_v9.py:20
    if rc == ERR_ARGUMENT or rc == ERR_WORKSPACE:
        raise ValueError(msg)
    raise B200Error(f"libfp8approx_b200 error {rc}: {msg}")
