"""Stage-by-stage check of the bf16 wide kernel set against the oracle (one layer, so every workspace
buffer still holds its stage's result after the step).  Run on the GPU box: python tools/debug_wide.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import torch.nn.functional as F
import __graft_entry__ as g
g.build()
import dlwp_benchmark_b200 as pkg
from oracle import wrappers as owrap
from helpers import synthetic

kw = dict(n_modes=[8, 8], constant_channels=2, prescribed_channels=1, prognostic_channels=3, hidden_channels=128,
          lifting_channels=256, projection_channels=256, n_layers=1, context_size=1)
B, H, W = 2, 12, 200
if len(sys.argv) > 1:
    kw["n_modes"] = [16, 64]; H, W = 21, 320
TW = (W + 127) // 128
torch.manual_seed(1234)
ref = owrap.FNO2DModule(**kw).eval()
mod = pkg.FNO2DModule(**kw).cuda().eval()
mod.load_state_dict(ref.state_dict(), strict=True)
mod.precision = pkg.native.PREC_BF16
const, presc, prog = synthetic(B, 2, 2, 1, 3, H, W)
with torch.no_grad():
    want = ref(constants=const, prescribed=presc, prognostic=prog)
    got = mod(constants=const.cuda(), prescribed=presc.cuda(), prognostic=prog.cuda()).cpu()
plan = list(mod._plans.values())[0]
nt = B * H * TW


def image(name, K):
    """[nt, 128 slots, K] float from the chunk-major bf16 tile images"""
    raw = plan.debug_read(name, nt * 128 * K * 2).view(torch.bfloat16).view(nt, K // 8, 128, 8)
    return raw.permute(0, 2, 1, 3).reshape(nt, 128, K).float().cpu()


def to_grid(img):
    """[nt, 128, K] -> [B, K, H, W]"""
    K = img.shape[-1]
    return img.view(B, H, TW * 128, K)[:, :, :W].permute(0, 3, 1, 2)


def rel(a, b):
    return float((a - b).abs().max() / b.abs().max())


fno = ref.fno
x_in = torch.cat([const[:, 0], presc[:, 0:1].flatten(1, 2), prog[:, 0:1].flatten(1, 2)], 1)
Cin = x_in.shape[1]
img_in = image("tcw_in", 32)
gin = to_grid(img_in)
print("gather hi+lo vs input:", rel(gin[:, :Cin] + gin[:, 16:16 + Cin], x_in), " ones:", float(gin[:, Cin].min()), float(gin[:, 15].min()))
with torch.no_grad():
    x0 = fno.lifting(x_in)
    print("lift (x0):", rel(to_grid(image("tcw_x0", 128)), x0))
    xf = torch.fft.rfft(x0, dim=-1, norm="forward")[..., :kw["n_modes"][1] // 2 + 1] / H   # [B, C, H, Kc]
    Kc = xf.shape[-1]
    HB = (H + 63) // 64
    sa = plan.debug_read("tcw_sa", B * Kc * HB * 32768).view(torch.bfloat16).view(B, Kc, HB, 16, 128, 8)
    sa = sa.permute(0, 1, 2, 4, 3, 5).reshape(B, Kc, HB * 64, 2, 128).float().cpu()[:, :, :H]      # [B, Kc, H, re|im, C]
    specA = torch.complex(sa[..., 0, :], sa[..., 1, :]).permute(0, 3, 2, 1)                       # [B, C, H, Kc]
    print("fwd W-DFT (spectrum images):", rel(torch.view_as_real(specA), torch.view_as_real(xf)))
    from oracle.neuralop_fno import neuralop_row_bins
    rin, rout = neuralop_row_bins(H, kw["n_modes"][0])
    hh = torch.arange(H, dtype=torch.float64)
    Fh = torch.exp(-2j * torch.pi * torch.tensor(rin, dtype=torch.float64)[:, None] * hh[None, :] / H)   # [R, H]
    X = torch.einsum("rh,bchk->bcrk", Fh, xf.to(torch.complex128))
    R = len(rin)
    specX = plan.debug_read("specX", R * Kc * B * 128 * 8).view(torch.float32).view(R, Kc, B, 128, 2).cpu()
    specX = torch.view_as_complex(specX.contiguous()).permute(2, 3, 0, 1)
    print("fwd H-DFT (specX):", rel(torch.view_as_real(specX), torch.view_as_real(X.to(torch.complex64))))
    y_spec = fno.fno_blocks.convs(x0, 0)
    x1 = y_spec + fno.fno_blocks.fno_skips[0](x0)
    print("layer (x1):", rel(to_grid(image("tcw_x1", 128)), x1), " [spectral part scale", float(y_spec.abs().max()), "]")
    incr = (want[:, 0] - prog[:, 0]).abs().max()
    print("output:", float((got[:, 0] - want[:, 0]).abs().max() / incr), "of increment", float(incr))
