"""GPU: dlwp_spectral_conv2d_static -- the allocation-free variant of the stand-alone Fourier layer whose scratch is
a file-scope device array (include/dlwp_b200.h) -- against the existing dlwp_spectral_conv2d entry point, through the C
ABI, on two calls in a row with different inputs and shapes (the array keeps the previous call's contents), and its
size limit.  Tolerance: the 2e-5 max-rel bound the other stand-alone spectral-layer tests use.

Collected last (file name) so that a failure here cannot hide other tests under ``pytest -x``.
"""
import ctypes as C

import pytest
import torch

from helpers import max_rel

pytestmark = pytest.mark.gpu

TOL = 2e-5


def _call(fn_name, x, weight, bias, H, W, rows, n_cols, Co, y=None):
    from dlwp_benchmark_b200 import native
    ri = native.int32_array(rows)
    y = torch.full((x.shape[0], Co, H, W), float("nan"), device="cuda") if y is None else y
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = getattr(native.lib(), fn_name)(
        C.c_void_p(x.data_ptr()), C.c_void_p(weight.data_ptr()), C.c_void_p(bias.data_ptr()) if bias is not None else None,
        C.c_void_p(y.data_ptr()), x.shape[0], x.shape[1], Co, H, W, len(rows), n_cols, ri, ri, native.FFT_NORM_BACKWARD, st)
    return rc, y


def _problem(seed, B, Cn, H, W, m1, m2, with_bias):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, Cn, H, W, generator=g).cuda()
    weight = torch.randn(Cn, Cn, 2 * m1, m2, 2, generator=g).cuda()
    bias = torch.randn(Cn, generator=g).cuda() if with_bias else None
    rows = list(range(m1)) + list(range(H - m1, H))
    return x, weight, bias, rows


def test_static_scratch_entry_matches_allocating_entry_on_two_calls_in_a_row(built_lib):
    from dlwp_benchmark_b200 import native
    assert torch.cuda.is_available()
    # warm-up call: the driver loads the kernels' code on their first launch, which takes device memory once
    x, weight, bias, rows = _problem(99, 1, 4, 8, 16, 2, 3, True)
    native.check(_call("dlwp_spectral_conv2d_static", x, weight, bias, 8, 16, rows, 3, 4)[0], "warm-up")
    torch.cuda.synchronize()
    # second problem: other shape, other data, smaller than the first (stale contents of the first call stay behind it)
    for seed, (B, Cn, H, W, m1, m2, with_bias) in enumerate([(3, 12, 32, 64, 6, 7, True), (2, 8, 16, 48, 4, 5, False)]):
        x, weight, bias, rows = _problem(100 + seed, B, Cn, H, W, m1, m2, with_bias)
        rc, want = _call("dlwp_spectral_conv2d", x, weight, bias, H, W, rows, m2, Cn)
        native.check(rc, "dlwp_spectral_conv2d")
        got = torch.full_like(want, float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc, got = _call("dlwp_spectral_conv2d_static", x, weight, bias, H, W, rows, m2, Cn, y=got)
        native.check(rc, "dlwp_spectral_conv2d_static")
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before, "the static-scratch entry took device memory"
        assert torch.isfinite(got).all()
        err = max_rel(got, want)
        print(f"call {seed}: static-scratch entry vs allocating entry, max-rel {err:.3e}")
        assert err < TOL


def test_static_scratch_entry_rejects_sizes_beyond_its_array(built_lib):
    from dlwp_benchmark_b200 import native
    x, weight, bias, rows = _problem(5, 64, 32, 64, 128, 6, 7, False)      # 64 MB of channels-last copies alone
    rc, _ = _call("dlwp_spectral_conv2d_static", x, weight, None, 64, 128, rows, 7, 32)
    assert rc == -1
    msg = native.lib().dlwp_last_error().decode()
    assert "bytes of scratch" in msg and str(native.STATIC_SCRATCH_BYTES) in msg
    # more retained rows than the by-value row table holds
    H = 2 * native.STATIC_MAX_ROWS + 8
    x, weight, bias, rows = _problem(6, 1, 2, H, 8, native.STATIC_MAX_ROWS // 2 + 1, 2, False)
    rc, _ = _call("dlwp_spectral_conv2d_static", x, weight, None, H, 8, rows, 2, 2)
    assert rc == -1 and "n_rows" in native.lib().dlwp_last_error().decode()
