"""Timestamps of fno_mid_kernel's passes (needs a DLWP_TRACE=1 build): block (0, 0), warp 1."""
import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
import bench
torch.manual_seed(0)
mod = pkg.TFNO2DModule(**bench.TFNO_C2).cuda().eval()
B, T, H, W = 256, 2, 32, 64
const, presc, prog = torch.randn(B,1,4,H,W,device='cuda'), torch.randn(B,T,1,H,W,device='cuda'), torch.randn(B,T,8,H,W,device='cuda')
tr = torch.zeros(64, dtype=torch.int64, device='cuda')
lib = native.lib()
with torch.no_grad():
    for _ in range(3): mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_mid(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
t = tr.cpu().view(8, 8)
t0 = int(t[0, 0])
print('prologue: start 0, before pdl_wait', int(t[0,1]) - t0, 'after pdl_wait', int(t[0,2]) - t0)
for p in range(1, 6):
    if t[p, 0] > 0:
        print('pass', p - 1, 'wait_start', int(t[p,0]) - t0, 'A_arrived', int(t[p,1]) - t0, 'phase1_done', int(t[p,2]) - t0,
              'phase2_done', int(t[p,3]) - t0, 'phase3_done', int(t[p,4]) - t0)
