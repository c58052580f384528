"""Debug aid: run the C2 model in fp32 and tensor-core mode on the same input and print errors."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import dlwp_benchmark_b200 as pkg
from dlwp_benchmark_b200 import native
from helpers import synthetic, max_rel
kw = dict(n_modes=[12, 12], constant_channels=4, prescribed_channels=1, prognostic_channels=8, hidden_channels=32,
          lifting_channels=256, projection_channels=256, n_layers=4, rank=0.8, context_size=1)
torch.manual_seed(1234)
a = pkg.TFNO2DModule(**kw).cuda().eval()
b = pkg.TFNO2DModule(**kw).cuda().eval()
b.load_state_dict(a.state_dict())
b.precision = native.PREC_FP32_TC
const, presc, prog = synthetic(8, 2, 4, 1, 8, 32, 64)
with torch.no_grad():
    ya = a(constants=const.cuda(), prescribed=presc.cuda(), prognostic=prog.cuda())
    yb = b(constants=const.cuda(), prescribed=presc.cuda(), prognostic=prog.cuda())
torch.cuda.synchronize()
print("tc used:", [p.uses_tensor_cores() for p in b._plans.values()])
print("max_rel tc vs fp32:", max_rel(yb, ya), "finite:", bool(torch.isfinite(yb).all()))
d = (yb - ya).abs()
print("err per channel:", d.amax(dim=(0, 1, 3, 4)).tolist())
print("err per row   :", d.amax(dim=(0, 1, 2, 4))[:8].tolist())
