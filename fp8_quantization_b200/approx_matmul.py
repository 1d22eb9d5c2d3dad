"""Host-side mirror of the reference's approx arithmetic module (same names, argument meaning and
error behaviour as approx/approx_matmul_whole_v9.py), backed by the sm_100a kernels.

Reference call edges replaced:
  custom_matmul_vectorize           approx/approx_matmul_whole_v9.py:10-169
  get_error_table_NN                :555-592
  float_to_fpany_absint_torch       :233-291
  fpany_absint_to_float_torch       :295-329
  quant_to_fp_any_vectorize_torch   :333-362
  param_prepare                     :189-229   (host scalars only; no kernel)
"""
import torch

from . import ops


def param_prepare(expo_width, mant_width, custom_bias=None, debug_mode=False):
    """Derived format constants (python-int branch of approx_matmul_whole_v9.py:189-229)."""
    fp_bias = custom_bias if custom_bias is not None else int((2 ** (expo_width - 1)) - 1)
    fp_bias = int(fp_bias)
    max_expo = int(2 ** expo_width - 1)
    return {
        "expo_width": expo_width, "mant_width": mant_width, "fp_bias": fp_bias, "bias_double": 2 * fp_bias,
        "max_norm": (2 ** (max_expo - fp_bias)) * (2 - 2 ** (-mant_width)),
        "min_norm": 2 ** (1 - fp_bias), "min_subnorm": (2 ** (1 - fp_bias)) * 2 ** (-mant_width),
        "max_expo": max_expo, "max_mant": int(2 ** mant_width - 1), "mant_scale": int(2 ** mant_width),
        "max_norm_int": int(2 ** (expo_width + mant_width) - 1), "OF_UF_mod": int(2 ** (expo_width + mant_width)),
    }


def get_error_table_NN(expo_width, mant_width, withComp, dnsmp_factor, device=None):
    """Same table choice and the same exceptions as the reference (:555-592).  Returned on `device`
    (default: current CUDA device, like the reference which builds its tables on 'cuda')."""
    t = ops.error_table(expo_width, mant_width, withComp, dnsmp_factor)
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
    return t.to(device)


def _is_tensor_bias(*biases):
    return any(torch.is_tensor(b) for b in biases)


def custom_matmul_vectorize(A, B, expo_width, mant_width, custom_bias_A, custom_bias_B, custom_bias_R,
                            error_table_NN, with_approx=True, with_s2nn2s_opt=False, sim_hw_add_OFUF=False,
                            with_OF_opt=False, with_UF_opt=False, golden_clip_OF=False, quant_btw_mult_accu=True,
                            debug_mode=False, self_check_mode=False):
    """[Mrows,K] x [K,N] -> [Mrows,N] with one bias per operand, as the reference function.

    Python-int biases select the regular numerics; int32 *tensor* biases select the reference's
    single-column quirk numerics (integer pow in param_prepare, approx_calculation.py:801-809).
    sim_hw_add_OFUF / with_OF_opt / with_UF_opt are ignored exactly as in v9.
    """
    assert A.shape[1] == B.shape[0]
    quirk = _is_tensor_bias(custom_bias_A, custom_bias_B, custom_bias_R)
    p = ops.make_params(dict(expo_width=expo_width, mant_width=mant_width, with_approx=with_approx,
                             with_s2nn2s_opt=with_s2nn2s_opt, golden_clip_OF=golden_clip_OF,
                             quant_btw_mult_accu=quant_btw_mult_accu), single_column_quirk=quirk,
                        allow_unpinned_format=True)
    to_int = lambda b: int(b.reshape(-1)[0].item()) if torch.is_tensor(b) else int(b)
    bB = custom_bias_B
    if torch.is_tensor(bB) and bB.numel() > 1:
        bB = bB.reshape(-1).to(torch.int32)
    else:
        bB = to_int(bB)
    table = error_table_NN if error_table_NN is not None else torch.zeros((2 ** mant_width,) * 2, dtype=torch.int8)
    return ops.approx_matmul(A, B, to_int(custom_bias_A), bB, to_int(custom_bias_R), p, custom_table=table)


def float_to_fpany_absint_torch(param_dict, values, clip_OF=False, return_extract=True):
    e, m = ops.fpany_encode(values, param_dict["expo_width"], param_dict["mant_width"], param_dict["fp_bias"],
                            clip_OF=clip_OF)
    if return_extract:
        return e, m
    return e * param_dict["mant_scale"] + m


def fpany_absint_to_float_torch(param_dict, sign=None, abs_int=None, expo=None, mant=None):
    if abs_int is not None:
        abs_int = torch.as_tensor(abs_int)
        expo = torch.div(abs_int, param_dict["mant_scale"], rounding_mode="floor")
        mant = abs_int % param_dict["mant_scale"]
    return ops.fpany_decode(expo, mant, param_dict["expo_width"], param_dict["mant_width"], param_dict["fp_bias"],
                            sign=sign)


def quant_to_fp_any_vectorize_torch(arr, expo_width, mant_width, custom_bias=None, clip_OF=True):
    pd = param_prepare(expo_width, mant_width, custom_bias)
    return ops.fpany_quant(arr, expo_width, mant_width, pd["fp_bias"], clip_OF=clip_OF)
