"""Timestamps of the lifting / projection MLP epilogue (DLWP_TRACE=1 build): block 0, worker 0, thread 0, per 16-column chunk step."""
import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
import bench
torch.manual_seed(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
which = sys.argv[2] if len(sys.argv) > 2 else 'proj'
mod = pkg.TFNO2DModule(**bench.TFNO_C2).cuda().eval()
T, H, W = 2, 32, 64
const, presc, prog = torch.randn(B,1,4,H,W,device='cuda'), torch.randn(B,T,1,H,W,device='cuda'), torch.randn(B,T,8,H,W,device='cuda')
tr = torch.zeros(64 * 8, dtype=torch.int64, device='cuda')
lib = native.lib()
with torch.no_grad():
    for _ in range(3): mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_mlp(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)      # lift writes first, proj overwrites: the last kernel's stamps stay
    torch.cuda.synchronize()
t = tr.cpu().view(64, 8)
t0 = int(t[8, 2])
print('chunk step: pass_wait_start d1_full | ld_start ld_done | compute_done a2_empty | stored arrived   (cycles; proj kernel, from chunk step 8)')
prev = t0
for c in range(8, 40):
    row = [int(x) - t0 if x > 0 else None for x in t[c]]
    print(c, row, 'step', (int(t[c, 7]) - prev))
    prev = int(t[c, 7])
