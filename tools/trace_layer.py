import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
import bench
kw = bench.TFNO_C2
torch.manual_seed(0)
mod = pkg.TFNO2DModule(**kw).cuda().eval()
B, T, H, W = 256, 2, 32, 64
const, presc, prog = torch.randn(B,1,4,H,W,device='cuda'), torch.randn(B,T,1,H,W,device='cuda'), torch.randn(B,T,8,H,W,device='cuda')
tr = torch.zeros(2*16*8, dtype=torch.int64, device='cuda')
lib = C.CDLL(native.LIB_PATH) if hasattr(native,'LIB_PATH') else native.lib()
with torch.no_grad():
    for _ in range(3): mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
t = tr.cpu().view(2,16,8)
names = ['load_issue','XFULL_seen','DEMPTY_seen','DFULL_epi','xt_arrive','xt_seen_fin','fin_done','d3_drained(prev)']
for w in range(2):
    t0 = int(t[w][t[w]>0].min())
    print('worker', w)
    for u in range(14):
        print(u, ' '.join(f'{names[e]}={int(t[w,u,e])-t0 if t[w,u,e]>0 else -1}' for e in range(8)))
