"""Timestamps of afno_mid_tc_kernel's stages (needs a DLWP_TRACE=1 build): block 0, control warp and epilogue warp 0."""
import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
import bench
torch.manual_seed(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
mod = pkg.FourCastNet(**bench.AFNO_C3).cuda().eval()
T, H, W = 2, 32, 64
const, presc, prog = torch.randn(B,1,4,H,W,device='cuda'), torch.randn(B,T,1,H,W,device='cuda'), torch.randn(B,T,3,H,W,device='cuda')
tr = torch.zeros(4 * 24, dtype=torch.int64, device='cuda')
lib = native.lib()
with torch.no_grad():
    for _ in range(3): mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_afno_mid(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
t = tr.cpu().view(4, 24)
t0 = int(t[0, 0])
names = {1: 'after_pdl_wait', 2: 'A_in', 3: 'G1_issued', 4: 'X_full', 5: 'G2a_issued', 6: 'H_full', 7: 'G2b_issued', 8: 'Y_full', 9: 'G3_issued',
         12: 'e:D1_full', 13: 'e:E1_done', 14: 'e:D2a_full', 15: 'e:E2a_done', 16: 'e:D2b_full', 17: 'e:E2b_done', 18: 'e:D3_full', 19: 'e:E3_done'}
print(f'B={B} (last block of the evaluation; cycles from kernel start)')
for it in range(4):
    ev = sorted((int(t[it, k]) - t0, names[k]) for k in names if t[it, k] > 0)
    if ev:
        print('item', it, ' '.join(f'{n}={c}' for c, n in ev))
