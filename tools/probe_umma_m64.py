import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo')
from dlwp_benchmark_b200 import native
M,N,K=64,32,32
torch.manual_seed(1)
a=torch.randn(M,K,device='cuda'); b=torch.randn(N,K,device='cuda')
d=torch.full((128,N),float('nan'),device='cuda')
st=C.c_void_p(torch.cuda.current_stream().cuda_stream)
native.check(native.lib().dlwp_selftest_umma(C.c_void_p(a.data_ptr()),C.c_void_p(b.data_ptr()),C.c_void_p(d.data_ptr()),M,N,K,st))
torch.cuda.synchronize()
want=a.bfloat16().float()@b.bfloat16().float().t()
# for each want row find matching lane
for m in range(0,64,8):
    diffs=(d[:, :N]-want[m][None,:]).abs().max(dim=1).values
    lane=int(diffs.argmin()); print('row',m,'-> lane',lane,'err',float(diffs[lane]))
