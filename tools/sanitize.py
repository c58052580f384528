"""compute-sanitizer driver: runs one case of the hand-written pipelines (mbarrier / TMEM / bulk-copy kernels in their
multi-tile steady state) without pytest, so that `compute-sanitizer --tool memcheck|racecheck|synccheck python
tools/sanitize.py <case>` instruments only this library's launches.  Cases:
  c4    TFNO2D of configs/model/fno.yaml, 32x64, `members` members (default 64), 2 lead steps, tcgen05 split-bf16 kernel set
  afno  FourCastNet AFNO (fourcastnet.yaml sizes), 24 members, 2 lead steps, tcgen05 AFNO kernel set
  wide  FNO2D at hidden width 128, bf16 tcgen05 kernel set, the `limits` grid of tests/test_tc_wide_gpu.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import dlwp_benchmark_b200 as pkg  # noqa: E402


def main():
    case = sys.argv[1]
    members = int(sys.argv[2]) if len(sys.argv) > 2 else None
    torch.manual_seed(0)
    if case == "c4":
        B = members or 64
        mod = pkg.TFNO2DModule(n_modes=[12, 12], constant_channels=4, prescribed_channels=1, prognostic_channels=8,
                               hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4, rank=0.8,
                               context_size=1).cuda().eval()
        shp = (B, 3, 4, 1, 8, 32, 64)
    elif case == "afno":
        B = members or 24
        mod = pkg.FourCastNet(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
                              prognostic_channels=3, embed_dim=64, depth=4, num_blocks=4, mlp_ratio=4.0, context_size=1,
                              use_pos_embed=True).cuda().eval()
        shp = (B, 3, 4, 1, 3, 32, 64)
    elif case == "wide":
        B = members or 3
        mod = pkg.FNO2DModule(n_modes=[4, 4], constant_channels=0, prescribed_channels=0, prognostic_channels=14,
                              hidden_channels=128, lifting_channels=256, projection_channels=256, n_layers=1,
                              context_size=1).cuda().eval()
        mod.precision = pkg.native.PREC_BF16
        shp = (B, 3, 0, 0, 14, 70, 256)
    else:
        raise SystemExit("unknown case " + case)
    B, T, Cc, Cp, Cg, H, W = shp
    const = torch.randn(B, 1, Cc, H, W, device="cuda") if Cc else None
    presc = torch.randn(B, T, Cp, H, W, device="cuda") if Cp else None
    prog = torch.randn(B, T, Cg, H, W, device="cuda")
    n0 = pkg.native.launch_count()
    with torch.no_grad():
        out = mod(constants=const, prescribed=presc, prognostic=prog)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    print(f"sanitize case {case}: B={B}, {pkg.native.launch_count() - n0} launches, output finite, "
          f"tensor cores: {[p.uses_tensor_cores() for p in mod._plans.values()]}")


if __name__ == "__main__":
    main()
