"""GPU: the tcgen05 building blocks (descriptor encodings, TMEM accumulator layout)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _umma(a, b, M, N, K):
    from dlwp_benchmark_b200 import native
    d = torch.full((128, N), float("nan"), device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    native.check(native.lib().dlwp_selftest_umma(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()),
                                                 C.c_void_p(d.data_ptr()), M, N, K, st))
    torch.cuda.synchronize()
    return d


@pytest.mark.parametrize("M,N,K", [(128, 32, 16), (128, 32, 64), (128, 256, 48), (128, 128, 32), (128, 16, 32),
                                   (64, 8, 64), (64, 16, 64), (128, 64, 32)])
def test_umma_m128_matches_bf16_matmul(built_lib, M, N, K):
    # (the M = 64 shapes with N = 8 / 16 are the per-mode products of fno_mid_tc_kernel)
    torch.manual_seed(M + N + K)
    a, b = torch.randn(M, K, device="cuda"), torch.randn(N, K, device="cuda")
    want = a.bfloat16().float() @ b.bfloat16().float().t()
    got = _umma(a.contiguous(), b.contiguous(), M, N, K)
    if M == 64:        # an M=64 accumulator keeps row r in TMEM lane 32 * (r // 16) + r % 16
        rows = torch.tensor([32 * (r // 16) + r % 16 for r in range(64)], device="cuda")
        got = got[rows]
    got = got[:M]
    err = float((got - want).abs().max())
    assert err < 1e-3, f"max abs err {err}"


def _umma_bmn(a, b_kn, M, N, K, variant):
    from dlwp_benchmark_b200 import native
    d = torch.full((128, N), float("nan"), device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    native.check(native.lib().dlwp_selftest_umma_bmn(C.c_void_p(a.data_ptr()), C.c_void_p(b_kn.data_ptr()),
                                                     C.c_void_p(d.data_ptr()), M, N, K, variant, st))
    torch.cuda.synchronize()
    return d


@pytest.mark.parametrize("M,N,K", [(64, 32, 128), (128, 32, 32), (64, 64, 64), (64, 256, 64), (64, 128, 128)])
def test_umma_mn_major_b_operand_reads_a_tile_image(built_lib, M, N, K):
    """The forward W-DFT feeds the activation tile image itself (pixels = K rows, channels contiguous)
    to tcgen05.mma as an MN-major B operand; this pins that descriptor encoding."""
    torch.manual_seed(M * N + K)
    a, b = torch.randn(M, K, device="cuda"), torch.randn(K, N, device="cuda")
    want = a.bfloat16().float() @ b.bfloat16().float()
    got = _umma_bmn(a.contiguous(), b.contiguous(), M, N, K, 1)
    if M == 64:        # an M=64 accumulator keeps row r in TMEM lane 32 * (r // 16) + r % 16
        rows = torch.tensor([32 * (r // 16) + r % 16 for r in range(64)], device="cuda")
        got = got[rows]
    err = float((got[:M] - want).abs().max())
    assert err < 2e-3, f"max abs err {err}"
