"""clock64 stamps of afno_fwd_tc_kernel's pipeline events (CTA 0, first 16 tiles).  Needs a trace build:
DLWP_TRACE=1 python dlwp-benchmark_b200/build.py --force"""
import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
import bench
kw = bench.AFNO_C3
torch.manual_seed(0)
mod = pkg.FourCastNet(**kw).cuda().eval()
B, T, H, W = 256, 2, 32, 64
const = torch.randn(B, 1, kw["constant_channels"], H, W, device='cuda')
presc = torch.randn(B, T, kw["prescribed_channels"], H, W, device='cuda')
prog = torch.randn(B, T, kw["prognostic_channels"], H, W, device='cuda')
tr = torch.zeros(16 * 8, dtype=torch.int64, device='cuda')
lib = native.lib()
with torch.no_grad():
    for _ in range(3):
        mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_afno(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
t = tr.cpu().view(16, 8)
names = ['W:xwait', 'W:xfull', 'W:yempty', 'W:yfull_arr', 'W:dfull(prev)', 'W:drained', 'C:yfull_seen', 'C:mma_issued']
t0 = int(t[t > 0].min())
for u in range(14):
    print(u, ' '.join(f'{names[e]}={int(t[u, e]) - t0 if t[u, e] > 0 else -1}' for e in range(8)))
