"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/fp8approx_b200.h declares; error codes map to the reference's exceptions; nothing computes
without a GPU (no CPU fallback)."""
import os
import re

import pytest
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "fp8approx_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported_and_bound():
    import fp8_quantization_b200 as pkg
    lib = pkg._lib.load()
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/fp8approx_b200.h but not exported"
        assert s in pkg._lib.SIGNATURES, f"{s} has no ctypes signature"
    assert set(pkg._lib.SIGNATURES) == set(syms)


def test_table_selector_matches_reference_rules():
    import numpy as np
    import fp8_quantization_b200 as pkg
    from oracle import oracle as orc
    for (E, M) in [(4, 3), (3, 4), (2, 5)]:
        for wc in (True, False):
            for dn in (3, 4, 5):
                assert np.array_equal(pkg.ops.error_table(E, M, wc, dn).numpy(), orc.get_error_table_NN(E, M, wc, dn))
    with pytest.raises(ValueError, match="Invalid combination"):
        pkg.ops.error_table(5, 2, True, 3)
    with pytest.raises(UnboundLocalError):
        pkg.ops.error_table(3, 4, True, 2)
    with pytest.raises(UnboundLocalError):
        pkg.ops.error_table(2, 5, True, 6)
    assert int(pkg.ops.error_table(5, 2, True, 3, allow_unpinned_format=True).abs().sum()) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    import fp8_quantization_b200 as pkg
    with pytest.raises(pkg._lib.B200Error):
        pkg.quantize_to_fp8_ste_MM(torch.randn(8), 8, torch.tensor([1.0]), torch.tensor([3.0]), 1)
    with pytest.raises(pkg._lib.B200Error):
        pkg.custom_matmul_vectorize(torch.zeros(2, 2), torch.zeros(2, 2), 4, 3, 8, 8, 8, None)
    import ctypes
    t = ctypes.c_double(0)
    assert pkg._lib.load().b200_int_peak_probe(10, ctypes.byref(t), None) == pkg._lib.ERR_NO_DEVICE


def test_param_struct_layout():
    import ctypes
    import fp8_quantization_b200 as pkg
    p = pkg.ops.make_params(dict(expo_width=3, mant_width=4, dnsmp_factor=5, withComp=False, golden_clip_OF=True))
    assert (p.expo_width, p.mant_width, p.dnsmp_factor, p.withComp, p.golden_clip_OF) == (3, 4, 5, 0, 1)
    # 3 x int32, 15 x uint8 (ofuf_mode, force_table16, prepared_b_ready last), padding, custom_table, prepared_b,
    # prepared_b_bytes -- static_assert'ed in csrc/runtime.cu
    assert ctypes.sizeof(pkg._lib.ApproxParams) == 56
    assert pkg._lib.ApproxParams.ofuf_mode.offset == 24
    assert pkg._lib.ApproxParams.prepared_b_ready.offset == 26
    assert pkg._lib.ApproxParams.custom_table.offset == 32
    assert pkg._lib.ApproxParams.prepared_b.offset == 40 and pkg._lib.ApproxParams.prepared_b_bytes.offset == 48
    assert pkg.ops.make_params(dict(ofuf_mode="v5")).ofuf_mode == 1 and pkg.ops.make_params({}).ofuf_mode == 0


def test_integration_doc_quotes_the_dropin_file():
    """INTEGRATION.md section B must show the file that tests/test_gpu_dropin.py installs on the reference classes."""
    import re
    root = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    m = re.search(r"<!-- dropin.py begin -->\n```python\n(.*?)```\n<!-- dropin.py end -->", doc, re.S)
    assert m, "INTEGRATION.md: dropin.py block not found"
    assert m.group(1) == open(os.path.join(root, "fp8_quantization_b200", "dropin.py")).read()
