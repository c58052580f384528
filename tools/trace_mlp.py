"""Timeline of the lifting kernel's MLP epilogue (block 0, worker 0, warp 0) from a DLWP_TRACE=1 build: clock64 stamps per
32-column chunk: 0 before the D1FULL wait, 1 after it, 3 accumulators in registers, 4 GELU + split done, 6 operand stores
complete, 7 A2FULL arrivals (0-3 and 6-7 once per 64-column pass).   DLWP_TRACE=1 python -c 'import __graft_entry__ as g; g.build()'"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import dlwp_benchmark_b200 as pkg  # noqa: E402
from dlwp_benchmark_b200 import native  # noqa: E402

kw = bench.TFNO_C2
torch.manual_seed(0)
mod = pkg.TFNO2DModule(**kw).cuda().eval()
B, T, H, W = 256, 2, 32, 64
const, presc, prog = torch.randn(B, 1, 4, H, W, device="cuda"), torch.randn(B, T, 1, H, W, device="cuda"), torch.randn(B, T, 8, H, W, device="cuda")
tr = torch.zeros(64 * 8, dtype=torch.int64, device="cuda")
lib = native.lib()
with torch.no_grad():
    for _ in range(3):
        mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_mlp(C.c_void_p(tr.data_ptr()))
    mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    lib.dlwp_debug_set_trace_mlp(C.c_void_p(0))
t = tr.cpu().view(64, 8)
t0 = int(t[t > 0].min())
prev_end = 0
for nc in range(0, 64, 2):
    e = [int(t[nc, k]) - t0 if t[nc, k] > 0 else -1 for k in range(8)]
    e1 = [int(t[nc + 1, k]) - t0 if t[nc + 1, k] > 0 else -1 for k in range(8)]
    print(f"pass {nc // 2:2d} (tile {nc // 8}, pass {(nc // 2) % 4}): gap since last arrive {e[0] - prev_end:6d}  wait D1FULL {e[1] - e[0]:5d}  "
          f"tmem ld {e[3] - e[1]:4d}  math0 {e[4] - e[3]:5d}  math1 {e1[4] - e[4]:5d}  st wait {e1[6] - e1[4]:4d}  arrive {e1[7] - e1[6]:4d}  "
          f"pass total {e1[7] - e[0]:5d}")
    prev_end = e1[7]
