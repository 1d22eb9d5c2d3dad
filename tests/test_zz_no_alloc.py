"""b200_qamaa_gemm_fbias_static: the allocation-free variant of the quantize_after_mult_and_add table GEMM (scratch in a
file-scope device array of the library instead of a caller workspace; no counterpart in the reference, added for
callers that cannot call cudaMalloc).  Called through the C ABI and compared bit for bit with the existing entry point
b200_qamaa_gemm_fbias, on two calls in a row of different shapes so that whatever the first call left in the shared
array must not leak into the second.  Collected last on purpose (pytest -x)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, Mrows, K, N, E, M):
    import fp8_quantization_b200 as pkg
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(Mrows, K, generator=g) * 0.7).cuda()
    w = (torch.randn(N, K, generator=g) * 0.05).cuda()
    mant = torch.tensor([float(M)])
    xq, bx = pkg.quantize_to_fp8_ste_MM(x, 8, x.abs().max().reshape(1), mant, 1)
    wq, bw = pkg.quantize_to_fp8_ste_MM(w, 8, w.abs().amax(dim=1), mant, 1)
    # the result quantizer's per-tensor range: a little below the largest exact product sum, so that clipping is live
    maxval = ((xq @ wq.t()).abs().max() * 0.75).reshape(1).float().contiguous()
    return xq.contiguous(), wq.contiguous(), bx.reshape(-1).float().contiguous(), bw.reshape(-1).float().contiguous(), maxval


def _run_both(lib, xq, wq, bx, bw, maxval, M):
    from fp8_quantization_b200 import _lib
    Mrows, K = xq.shape
    N = wq.shape[0]
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    y = wq.t()                                             # [K, N] view: sb_k = 1, sb_n = K
    need = lib.b200_qamaa_gemm_workspace_bytes(Mrows, K, N)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    want = torch.full((Mrows, N), float("nan"), device="cuda")
    got = torch.full((Mrows, N), float("nan"), device="cuda")
    _lib.check(lib.b200_qamaa_gemm_fbias(p(xq), xq.stride(0), p(y), y.stride(0), y.stride(1), p(want), want.stride(0),
                                         Mrows, K, N, p(maxval), 8, float(M), 1, p(bx), p(bw), 1, p(ws), need, st))
    path_want = lib.b200_last_path().decode()
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(lib.b200_qamaa_gemm_fbias_static(p(xq), xq.stride(0), p(y), y.stride(0), y.stride(1), p(got),
                                                got.stride(0), Mrows, K, N, p(maxval), 8, float(M), 1, p(bx), p(bw), 1, st))
    path_got = lib.b200_last_path().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return want, got, path_want, path_got


def test_static_scratch_entry_point_matches_workspace_entry_point_twice_in_a_row():
    import fp8_quantization_b200 as pkg
    lib = pkg._lib.load()
    assert lib.b200_static_scratch_bytes() == 64 << 20
    # two calls in a row; the second is smaller and of another format, so stale planes / table / partial sums of the
    # first call lie beyond and beneath what it writes
    for seed, (Mrows, K, N, E, M) in enumerate([(1000, 520, 392, 4, 3), (257, 200, 72, 3, 4), (1000, 520, 392, 4, 3)]):
        xq, wq, bx, bw, maxval = _case(100 + seed, Mrows, K, N, E, M)
        want, got, path_want, path_got = _run_both(lib, xq, wq, bx, bw, maxval, M)
        assert path_got == path_want and path_got.startswith("qamaa-lut"), (path_want, path_got)
        assert not torch.isnan(got).any()
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), \
            f"static-scratch entry point differs from b200_qamaa_gemm_fbias ({Mrows}x{K}x{N} E{E}M{M})"
        # and both are the function itself: the arithmetic kernel (no table, no scratch)
        ref = pkg.ops.qamaa_matmul(xq, wq.t(), 8, maxval, float(M), 1)
        assert torch.equal(got.view(torch.int32), ref.view(torch.int32))


def test_static_scratch_entry_point_rejects_oversized_problems():
    import fp8_quantization_b200 as pkg
    lib = pkg._lib.load()
    Mrows = K = N = 4096
    assert lib.b200_qamaa_gemm_workspace_bytes(Mrows, K, N) > lib.b200_static_scratch_bytes()
    one = torch.ones(1, device="cuda")
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    # sizes are checked before any pointer is touched
    rc = lib.b200_qamaa_gemm_fbias_static(p(one), K, p(one), 1, K, p(one), N, Mrows, K, N, p(one), 8, 3.0, 1, p(one), p(one),
                                          0, None)
    assert rc == pkg._lib.ERR_WORKSPACE
    msg = lib.b200_last_error().decode()
    assert "static" in msg and str(64 << 20) in msg, msg
