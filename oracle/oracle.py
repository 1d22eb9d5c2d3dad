"""ctypes front end of the CPU oracle (oracle/approx_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
Parity: pinned against the unmodified reference through tests/golden (see approx_oracle.c).
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

i64 = ctypes.c_int64
c_int = ctypes.c_int
c_float = ctypes.c_float


class OrcCfg(ctypes.Structure):
    _fields_ = [("expo_width", c_int), ("mant_width", c_int), ("with_approx", c_int),
                ("with_s2nn2s_opt", c_int), ("golden_clip_OF", c_int),
                ("quant_btw_mult_accu", c_int), ("tensor_bias_quirk", c_int), ("sum_mode", c_int)]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "approx_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.orc_fp8_bias.restype = c_float
        _LIB.orc_fp8_bias.argtypes = [c_float, c_int, c_float, c_int]
    return _LIB


def _p(a, t=ctypes.c_void_p):
    return a.ctypes.data_as(t) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


_TABLES = None


def comp_tables():
    """The six tables of approx_matmul_whole_v9.py:393-549 (fixture extracted by make_tables.py)."""
    global _TABLES
    if _TABLES is None:
        z = np.load(os.path.join(_HERE, "..", "tests", "golden", "comp_tables.npz"))
        _TABLES = [z[f"t{i}"].astype(np.int8) for i in range(6)]
    return _TABLES


def get_error_table_NN(expo_width, mant_width, withComp, dnsmp_factor):
    """Table choice, approx/approx_matmul_whole_v9.py:555-592 (errors mirrored)."""
    L = comp_tables()
    n = 2 ** mant_width
    zeros = np.zeros((n, n), dtype=np.int8)
    if (expo_width, mant_width) == (4, 3):
        return zeros if withComp else L[0]
    if (expo_width, mant_width) == (3, 4):
        if withComp:
            if dnsmp_factor == 3:
                return L[1]
            if dnsmp_factor >= 4:
                return zeros
            raise UnboundLocalError("error_table_NN")
        return L[2]
    if (expo_width, mant_width) == (2, 5):
        if withComp:
            if dnsmp_factor in (3, 4):
                return L[3] if dnsmp_factor == 3 else L[4]
            if dnsmp_factor == 5:
                return zeros
            raise UnboundLocalError("error_table_NN")
        return L[5]
    raise ValueError("Invalid combination of expo_width and mant_width")


def make_cfg(E, M, with_approx=True, with_s2nn2s_opt=False, golden_clip_OF=False,
             quant_btw_mult_accu=True, tensor_bias_quirk=False, sum_mode=0):
    return OrcCfg(E, M, int(with_approx), int(with_s2nn2s_opt), int(golden_clip_OF),
                  int(quant_btw_mult_accu), int(tensor_bias_quirk), int(sum_mode))


def encode(v, E, M, bias, quirk=False, clip_OF=False):
    v = _f32(v)
    e = np.empty(v.shape, np.int32)
    m = np.empty(v.shape, np.int32)
    lib().orc_encode_array(_p(v), i64(v.size), E, M, int(bias), int(quirk), int(clip_OF), _p(e), _p(m))
    return e, m


def decode(expo, mant, E, M, bias, sign=None):
    expo = np.ascontiguousarray(expo, np.int32)
    mant = np.ascontiguousarray(mant, np.int32)
    sign = _f32(sign) if sign is not None else None
    out = np.empty(expo.shape, np.float32)
    lib().orc_decode_array(_p(expo), _p(mant), _p(sign), i64(expo.size), E, M, int(bias), _p(out))
    return out


def quant(v, E, M, bias, quirk=False, clip_OF=True):
    v = _f32(v)
    out = np.empty_like(v)
    lib().orc_quant_array(_p(v), i64(v.size), E, M, int(bias), int(quirk), int(clip_OF), _p(out))
    return out


def pair(a, b, table, bA, bB, bR, **cfg_kw):
    """Addend of one operand pair (the K=1 view of custom_matmul_vectorize)."""
    a, b = np.broadcast_arrays(_f32(a), _f32(b))
    a, b = _f32(a), _f32(b)
    cfg = make_cfg(**cfg_kw)
    T = np.ascontiguousarray(table, np.int8)
    out = np.empty(a.shape, np.float32)
    lib().orc_pair_array(ctypes.byref(cfg), _p(T), _p(a), _p(b), i64(a.size), int(bA), int(bB), int(bR),
                         _p(out))
    return out


def approx_matmul(A, B, table, bA, bB, bR, n_begin=0, n_end=0, **cfg_kw):
    """C = approx_multiply(A [Mrows,K], B [K,N]) with per-column weight bias bB (int or [N])."""
    A = _f32(A)
    B = _f32(B)
    Mr, K = A.shape
    K2, N = B.shape
    assert K == K2
    cfg = make_cfg(**cfg_kw)
    T = np.ascontiguousarray(table, np.int8)
    bBv = np.ascontiguousarray(np.broadcast_to(np.asarray(bB, np.int32).reshape(-1), (N,)), np.int32)
    C = np.zeros((Mr, N), np.float32)
    lib().orc_approx_matmul(ctypes.byref(cfg), _p(T), _p(A), i64(K), i64(1), _p(B), i64(N), i64(1),
                            _p(C), i64(N), i64(1), i64(Mr), i64(K), i64(N),
                            int(bA), _p(bBv), 1, int(bR), i64(n_begin), i64(n_end))
    if n_end:
        return C[:, n_begin:n_end]
    return C


def approx_matmul_abs(A, B, table, bA, bB, bR, **cfg_kw):
    A = _f32(A)
    B = _f32(B)
    Mr, K = A.shape
    N = B.shape[1]
    cfg = make_cfg(**cfg_kw)
    T = np.ascontiguousarray(table, np.int8)
    bBv = np.ascontiguousarray(np.broadcast_to(np.asarray(bB, np.int32).reshape(-1), (N,)), np.int32)
    C = np.zeros((Mr, N), np.float64)
    lib().orc_approx_matmul_abs(ctypes.byref(cfg), _p(T), _p(A), i64(K), i64(1), _p(B), i64(N), i64(1),
                                _p(C), i64(Mr), i64(K), i64(N), int(bA), _p(bBv), 1, int(bR))
    return C


def fp8_bias(maxval, n_bits=8, mantissa_bits=3.0, sign_bits=1):
    return lib().orc_fp8_bias(c_float(maxval), n_bits, c_float(mantissa_bits), sign_bits)


def fp8_fakequant(x, maxval, n_bits=8, mantissa_bits=3.0, sign_bits=1):
    """quantize_to_fp8_ste_MM; maxval scalar/[1] (per tensor) or [C] (per dim-0 channel)."""
    x = _f32(x)
    mv = _f32(np.asarray(maxval).reshape(-1))
    C = mv.size
    inner = x.size // x.shape[0] if (C > 1 and x.ndim > 0) else 1
    if C > 1:
        assert x.shape[0] == C
    y = np.empty_like(x)
    bias = np.empty(C, np.float32)
    lib().orc_fp8_fakequant(_p(x), _p(y), i64(x.size), _p(mv), i64(C), i64(inner), n_bits,
                            c_float(mantissa_bits), sign_bits, _p(bias))
    return y, bias


def qamaa_matmul(A, B, maxval, n_bits=8, mantissa_bits=3.0, sign_bits=1, sum_mode=0):
    A = _f32(A)
    B = _f32(B)
    Mr, K = A.shape
    N = B.shape[1]
    C = np.zeros((Mr, N), np.float32)
    lib().orc_qamaa_matmul(_p(A), i64(K), i64(1), _p(B), i64(N), i64(1), _p(C), i64(N), i64(1),
                           i64(Mr), i64(K), i64(N), c_float(maxval), n_bits, c_float(mantissa_bits),
                           sign_bits, sum_mode)
    return C


def allnorm_encode(v, E, M, bias, clip_OF=False):
    v = _f32(v)
    e = np.empty(v.shape, np.int32)
    m = np.empty(v.shape, np.int32)
    lib().orc_allnorm_encode_array(_p(v), i64(v.size), E, M, int(bias), int(clip_OF), _p(e), _p(m))
    return e, m


def allnorm_decode(expo, mant, M, bias, sign=None):
    expo = np.ascontiguousarray(expo, np.int32)
    mant = np.ascontiguousarray(mant, np.int32)
    sign = _f32(sign) if sign is not None else None
    out = np.empty(expo.shape, np.float32)
    lib().orc_allnorm_decode_array(_p(expo), _p(mant), _p(sign), i64(expo.size), M, int(bias), _p(out))
    return out


def v5_ofuf(x_int, y_int, comp, B_neg, E, M, sim_hw_add_OFUF, with_OF_opt, with_UF_opt):
    x_int = np.ascontiguousarray(x_int, np.int32)
    y_int = np.ascontiguousarray(y_int, np.int32)
    comp = np.ascontiguousarray(comp, np.int32)
    out = np.empty(x_int.shape, np.int32)
    lib().orc_v5_ofuf_array(_p(x_int), _p(y_int), _p(comp), i64(x_int.size), int(B_neg), E, M,
                            int(sim_hw_add_OFUF), int(with_OF_opt), int(with_UF_opt), _p(out))
    return out


_V5_TABLES = None


def v5_comp_table(E, M, withComp, dnsmp_factor):
    """get_comp_table_NN of approx/approx_matmul_whole_v5.py:516-542; without compensation the reference assigns an
    all-zero table and then raises ValueError -- the zero table is what the opt-in mode uses there."""
    global _V5_TABLES
    if _V5_TABLES is None:
        z = np.load(os.path.join(_HERE, "..", "tests", "golden", "comp_tables_v5.npz"))
        _V5_TABLES = [z[f"t{i}"].astype(np.int32) for i in range(6)]
    L = _V5_TABLES
    if not withComp:
        return np.zeros((2 ** M, 2 ** M), np.int32)
    if (E, M) == (4, 3):
        return L[0]
    if (E, M) == (3, 4):
        if dnsmp_factor == 3:
            return L[1]
        if dnsmp_factor >= 4:
            return L[2]
        raise UnboundLocalError("comp_table_NN")
    if (E, M) == (2, 5):
        if dnsmp_factor in (3, 4):
            return L[dnsmp_factor]
        if dnsmp_factor >= 5:
            return L[5]
        raise UnboundLocalError("comp_table_NN")
    raise ValueError("Invalid combination of expo_width and mant_width")


def v5_matmul(A, B, comp, E, M, bA, bB, bR, sim_hw_add_OFUF=False, with_OF_opt=False, with_UF_opt=False, sum_mode=0):
    """custom_matmul_vectorize of approx/approx_matmul_whole_v5.py:10-152 (see orc_v5_matmul)."""
    A = _f32(A)
    B = _f32(B)
    Mr, K = A.shape
    N = B.shape[1]
    comp = np.ascontiguousarray(comp, np.int32) if comp is not None else None
    bBv = np.ascontiguousarray(np.broadcast_to(np.asarray(bB, np.int32).reshape(-1), (N,)), np.int32)
    C = np.zeros((Mr, N), np.float32)
    lib().orc_v5_matmul(int(E), int(M), _p(comp), _p(A), i64(K), i64(1), _p(B), i64(N), i64(1), _p(C), i64(N), i64(1),
                        i64(Mr), i64(K), i64(N), int(bA), _p(bBv), 1, int(bR), int(sim_hw_add_OFUF), int(with_OF_opt),
                        int(with_UF_opt), int(sum_mode))
    return C


def im2col(x, kh, kw, stride, padding, dilation):
    x = _f32(x)
    B, C, H, W = x.shape
    Ho = (H + 2 * padding[0] - dilation[0] * (kh - 1) - 1) // stride[0] + 1
    Wo = (W + 2 * padding[1] - dilation[1] * (kw - 1) - 1) // stride[1] + 1
    col = np.empty((B * Ho * Wo, C * kh * kw), np.float32)
    lib().orc_im2col(_p(x), B, C, H, W, kh, kw, stride[0], stride[1], padding[0], padding[1],
                     dilation[0], dilation[1], _p(col))
    return col, Ho, Wo


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    """OpenMP thread count of the port, whatever OMP_NUM_THREADS says (torchrun exports 1 to its workers)."""
    lib().orc_set_num_threads(int(n))
    return num_threads()
