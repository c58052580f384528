import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
from dlwp_benchmark_b200 import native
for variant in (1, 2):
    for (M, N, K) in [(128, 32, 32), (64, 32, 128)]:
        torch.manual_seed(1)
        a = torch.randn(M, K, device='cuda'); b = torch.randn(K, N, device='cuda')
        d = torch.full((128, N), float('nan'), device='cuda')
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        native.check(native.lib().dlwp_selftest_umma_bmn(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(d.data_ptr()), M, N, K, variant, st))
        torch.cuda.synchronize()
        want = a.bfloat16().float() @ b.bfloat16().float()
        if M == 64:
            rows = torch.tensor([32 * (r // 16) + r % 16 for r in range(64)], device='cuda')
            d = d[rows]
        print('variant', variant, (M, N, K), 'max err', float((d[:M] - want).abs().max()))
