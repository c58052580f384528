"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the committed
golden fixtures.  Tolerance for fp32: max|y - y_ref| / max|y_ref| <= 1e-4 on a single step
(BASELINE.json north_star); rollout RMSE curves within 1e-3 relative of the oracle's curve.

FNO/TFNO parity is against a restatement; the reference's neuralop dependency is not available.
AFNO and SpectralConv2d parity is pinned by fixtures the unmodified reference source produced.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from helpers import golden_state_dict, load_golden, max_rel, synthetic

pytestmark = pytest.mark.gpu

TOL_STEP = 1e-4
TOL_CURVE = 1e-3


@pytest.fixture(scope="module")
def pkg(built_lib):
    import dlwp_benchmark_b200 as p
    assert torch.cuda.is_available()
    return p


def _dev(t):
    return None if t is None else t.cuda()


def _rmse_curve(out, tgt):
    from oracle import metrics as om
    r, _ = om.rmse_acc(out.double(), tgt.double())
    return r


# ---------------------------------------------------------------------------------------
# stand-alone spectral layer (unet.py:19-69) against the reference-generated fixture
# ---------------------------------------------------------------------------------------
def test_spectral_conv2d_matches_reference_golden(pkg):
    from dlwp_benchmark_b200 import native
    g = load_golden("spectral2d")
    x0 = torch.from_numpy(g["x"])
    B, Ci, H, W = x0.shape
    m1, m2 = 6, 5
    w1 = torch.from_numpy(g["conv/weights1"])          # [Ci, Co, m1, m2, 2]
    w2 = torch.from_numpy(g["conv/weights2"])
    Co = w1.shape[1]
    # the stand-alone entry wants Ci == Co: zero-pad the input channels (exact)
    x = torch.zeros(B, Co, H, W)
    x[:, :Ci] = x0
    weight = torch.zeros(Co, Co, 2 * m1, m2, 2)
    weight[:Ci, :, :m1] = w1                           # rows 0..m1-1
    weight[:Ci, :, m1:] = w2                           # rows H-m1..H-1
    assert max_rel(_spectral(x.cuda(), weight.cuda(), H, W, m1, m2), torch.from_numpy(g["y_conv"])) < 2e-5


def test_spectral_conv2d_module_and_residual_block_match_reference_golden(pkg):
    """The drop-in modules (reference parameter names; in_channels != out_channels) against the fixtures the unmodified
    reference produced: SpectralConv2d 8 -> 12, FourierResidualBlock 8 -> 12 with the shortcut convolution, and a
    GroupNorm block (norm=True, two groups, identity shortcut)."""
    g = load_golden("spectral2d")
    x = torch.from_numpy(g["x"]).cuda()
    conv = pkg.SpectralConv2d(8, 12, 6, 5).cuda()
    conv.load_state_dict({k[5:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("conv/")}, strict=True)
    blk = pkg.FourierResidualBlock(8, 12, modes1=6, modes2=5).cuda()
    blk.load_state_dict({k[6:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("block/")}, strict=True)
    with torch.no_grad():
        assert max_rel(conv(x), torch.from_numpy(g["y_conv"])) < 2e-5
        for _ in range(2):      # the second call reuses the module's workspace
            assert max_rel(blk(x), torch.from_numpy(g["y_block"])) < 2e-5
    g = load_golden("spectral2d_norm")
    nblk = pkg.FourierResidualBlock(12, 12, modes1=5, modes2=7, norm=True, n_groups=2).cuda()
    nblk.load_state_dict({k[6:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("block/")}, strict=True)
    with torch.no_grad():
        err = max_rel(nblk(torch.from_numpy(g["x"]).cuda()), torch.from_numpy(g["y_block"]))
    print(f"FourierResidualBlock with GroupNorm vs reference fixture: max-rel {err:.3e}")
    assert err < 2e-5
    with pytest.raises(NotImplementedError):
        blk(x)      # grad mode on: forward-only


def _spectral(x, weight, H, W, m1, m2):
    from dlwp_benchmark_b200 import native
    rows = list(range(m1)) + list(range(H - m1, H))
    ri = native.int32_array(rows)
    Cn = weight.shape[0]
    y = torch.empty(x.shape[0], Cn, H, W, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    native.check(native.lib().dlwp_spectral_conv2d(
        C.c_void_p(x.data_ptr()), C.c_void_p(weight.data_ptr()), None, C.c_void_p(y.data_ptr()),
        x.shape[0], Cn, Cn, H, W, 2 * m1, m2, ri, ri, native.FFT_NORM_BACKWARD, st))
    torch.cuda.synchronize()
    return y


def test_spectral_layer_is_linear_and_shift_equivariant(pkg):
    torch.manual_seed(0)
    B, Cn, H, W, m1, m2 = 2, 6, 16, 32, 4, 5
    weight = torch.randn(Cn, Cn, 2 * m1, m2, 2, device="cuda")
    conv = lambda x: _spectral(x.contiguous(), weight, H, W, m1, m2)
    a, b = torch.randn(B, Cn, H, W, device="cuda"), torch.randn(B, Cn, H, W, device="cuda")
    assert max_rel(conv(2.0 * a - 3.0 * b), 2.0 * conv(a) - 3.0 * conv(b)) < 2e-5
    assert max_rel(conv(torch.roll(a, 5, dims=-1)), torch.roll(conv(a), 5, dims=-1)) < 2e-5
    assert max_rel(conv(torch.roll(a, 3, dims=-2)), torch.roll(conv(a), 3, dims=-2)) < 2e-5


# ---------------------------------------------------------------------------------------
# FNO2D / TFNO2D (BASELINE configs C1, C1b, C2) against the oracle restatement
# ---------------------------------------------------------------------------------------
FNO_CASES = {
    "C1": ("FNO2DModule", dict(n_modes=[12, 12], constant_channels=0, prescribed_channels=0, prognostic_channels=1,
                               hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4,
                               context_size=1), (8, 64, 64)),
    "C1b": ("FNO2DModule", dict(n_modes=[12, 12], constant_channels=0, prescribed_channels=0, prognostic_channels=1,
                                hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4,
                                context_size=10), (4, 64, 64)),
    "C2": ("TFNO2DModule", dict(n_modes=[12, 12], constant_channels=4, prescribed_channels=1, prognostic_channels=8,
                                hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4,
                                rank=0.8, bias=True, context_size=1, type="TFNO2DModule", name="tfno"), (8, 32, 64)),
    "odd": ("FNO2DModule", dict(n_modes=[6, 10], constant_channels=1, prescribed_channels=0, prognostic_channels=2,
                                hidden_channels=12, lifting_channels=24, projection_channels=20, n_layers=2,
                                context_size=2), (3, 17, 30)),
}


def _fno_pair(pkg, case):
    from oracle import wrappers as owrap
    cls, kw, (B, H, W) = FNO_CASES[case]
    torch.manual_seed(1234)
    ref = getattr(owrap, cls)(**kw).eval()
    mod = getattr(pkg, cls)(**kw).cuda().eval()
    mod.load_state_dict(ref.state_dict(), strict=True)
    mod.precision = pkg.native.PREC_FP32          # the fp32 FMA kernel set; the tensor-core tests override this
    return ref, mod, kw, (B, H, W)


@pytest.mark.parametrize("case", list(FNO_CASES))
def test_fno_single_step_matches_oracle(pkg, case):
    ref, mod, kw, (B, H, W) = _fno_pair(pkg, case)
    ctx = kw["context_size"]
    const, presc, prog = synthetic(B, ctx + 1, kw["constant_channels"], kw["prescribed_channels"],
                                   kw["prognostic_channels"], H, W)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog))
    assert got.is_cuda and got.shape == want.shape
    assert max_rel(got, want) < TOL_STEP
    # against fp64 evaluation of the same oracle (ground truth)
    ref64 = ref.double()
    with torch.no_grad():
        want64 = ref64(constants=None if const is None else const.double(),
                       prescribed=None if presc is None else presc.double(), prognostic=prog.double())
    assert max_rel(got, want64) < TOL_STEP


@pytest.mark.parametrize("case,steps", [("C1", 50), ("C1b", 12), ("C2", 20), ("odd", 5)])
def test_fno_rollout_tracks_oracle_rmse_curve(pkg, case, steps):
    ref, mod, kw, (B, H, W) = _fno_pair(pkg, case)
    ctx = kw["context_size"]
    T = ctx + steps
    const, presc, prog = synthetic(B, T, kw["constant_channels"], kw["prescribed_channels"],
                                   kw["prognostic_channels"], H, W, smooth=True)
    tgt = synthetic(B, steps, 0, 0, kw["prognostic_channels"], H, W, seed=77)[2]
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
    assert got.shape == (B, steps, kw["prognostic_channels"], H, W)
    assert torch.isfinite(got).all()
    rw, rg = _rmse_curve(want, tgt), _rmse_curve(got, tgt)
    assert float(((rg - rw).abs() / rw).max()) < TOL_CURVE
    assert max_rel(got[:, 0], want[:, 0]) < TOL_STEP


# tensor-core mode (DLWP_PREC_FP32_TC): split-bf16 products on tcgen05, same 1e-4 bound
@pytest.mark.parametrize("case,steps", [("C2", 1), ("C1", 1), ("C2", 20), ("C1", 50), ("C1b", 12)])
def test_fno_tensor_core_path_matches_oracle(pkg, case, steps):
    # (C1b: context_size 10 -- ten input channels gathered from ground-truth frames and the model's own outputs, fno.py:89-100)
    from dlwp_benchmark_b200 import native
    ref, mod, kw, (B, H, W) = _fno_pair(pkg, case)
    mod.precision = native.PREC_FP32_TC
    ctx = kw["context_size"]
    const, presc, prog = synthetic(B, ctx + steps, kw["constant_channels"], kw["prescribed_channels"],
                                   kw["prognostic_channels"], H, W, smooth=steps > 1)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
    assert all(p.uses_tensor_cores() for p in mod._plans.values()), "tcgen05 path was not taken"
    assert torch.isfinite(got).all()
    assert max_rel(got[:, 0], want[:, 0]) < TOL_STEP
    if steps > 1:
        tgt = synthetic(B, steps, 0, 0, kw["prognostic_channels"], H, W, seed=77)[2]
        rw, rg = _rmse_curve(want, tgt), _rmse_curve(got, tgt)
        assert float(((rg - rw).abs() / rw).max()) < TOL_CURVE


def test_tensor_core_mode_falls_back_to_fp32_kernels_on_other_shapes(pkg):
    from dlwp_benchmark_b200 import native
    ref, mod, kw, (B, H, W) = _fno_pair(pkg, "odd")
    mod.precision = native.PREC_FP32_TC
    const, presc, prog = synthetic(B, 3, 1, 0, 2, H, W)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=None, prognostic=_dev(prog))
    assert not any(p.uses_tensor_cores() for p in mod._plans.values())
    assert max_rel(got, want) < TOL_STEP


def test_cpu_inputs_and_grad_mode_guards(pkg):
    _, mod, kw, (B, H, W) = _fno_pair(pkg, "odd")
    const, presc, prog = synthetic(B, 3, 1, 0, 2, H, W)
    with torch.no_grad(), pytest.raises(RuntimeError):
        mod(constants=const, prescribed=None, prognostic=prog)
    with pytest.raises(NotImplementedError):      # the host entry is inference-only
        mod.forward_host(constants=const, prescribed=None, prognostic=prog)
    # grad mode returns a tensor with a graph (training path, tests/test_training_gpu.py), never one without
    amod = pkg.FourCastNet(img_height=16, img_width=32, patch_size=[1, 1], constant_channels=0, prescribed_channels=0,
                           prognostic_channels=3, embed_dim=32, depth=1, num_blocks=4, context_size=1).cuda()
    assert amod(prognostic=torch.randn(1, 2, 3, 16, 32, device="cuda")).requires_grad
    with pytest.raises(NotImplementedError):
        amod.forward_host(prognostic=torch.randn(1, 2, 3, 16, 32))


# ---------------------------------------------------------------------------------------
# FourCastNet (C3) against reference-generated fixtures and the oracle
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["afno_c3", "afno_small", "afno_patch"])
def test_afno_matches_reference_golden(pkg, name):
    g = load_golden(name)
    mod = pkg.FourCastNet(**g["kwargs"]).cuda().eval()
    mod.load_state_dict(golden_state_dict(g), strict=True)
    t = lambda k: torch.from_numpy(g[k]).cuda()
    ctx = g["kwargs"]["context_size"]
    with torch.no_grad():
        one = mod(constants=t("constants"), prescribed=t("prescribed")[:, :ctx + 1],
                  prognostic=t("prognostic")[:, :ctx + 1])
        roll = mod(constants=t("constants"), prescribed=t("prescribed"), prognostic=t("prognostic"))
    assert max_rel(one, torch.from_numpy(g["one_step"])) < TOL_STEP
    assert max_rel(roll, torch.from_numpy(g["rollout"])) < 2 * TOL_STEP


def test_afno_paper_size_matches_oracle(pkg):
    from oracle import afno as oafno
    kw = dict(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
              prognostic_channels=3, embed_dim=272, depth=2, num_blocks=8, context_size=1, use_pos_embed=False)
    torch.manual_seed(5)
    ref = oafno.AFNONet(**kw).eval()
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if "filter" in n:
                p.mul_(6.0)
    mod = pkg.FourCastNet(**kw).cuda().eval()
    mod.load_state_dict(ref.state_dict(), strict=True)
    const, presc, prog = synthetic(2, 3, 4, 1, 3, 32, 64)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog))
    assert max_rel(got, want) < TOL_STEP


@pytest.mark.parametrize("B,nb,frac,H,W,Cg", [(3, 4, 1.0, 32, 64, 3), (24, 4, 1.0, 32, 64, 3), (5, 8, 0.5, 16, 32, 4),
                                               (2, 4, 0.7, 24, 48, 6),
                                               # afno_mid_tc_kernel: 8 channel blocks of 8; truncated modes (16 rows x 8 columns); one sample
                                               (4, 8, 1.0, 32, 64, 3), (2, 4, 0.5, 32, 64, 3), (1, 4, 1.0, 32, 64, 3)])
def test_afno_tensor_core_kernels_match_oracle_and_fp32_kernel_set(pkg, B, nb, frac, H, W, Cg):
    """configs/model/fourcastnet.yaml shapes (embed_dim 64, mlp_ratio 4): the tcgen05 AFNO kernels of the
    default precision mode against the oracle and against the fp32 FMA kernel set; B = 24 gives every
    persistent CTA several 128-token tiles; more than 4 prognostic channels keep the head a separate launch."""
    from oracle import afno as oafno
    native = pkg.native
    kw = dict(img_height=H, img_width=W, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
              prognostic_channels=Cg, embed_dim=64, depth=4, num_blocks=nb, mlp_ratio=4.0, context_size=1,
              use_pos_embed=(Cg != 4), hard_thresholding_fraction=frac)
    torch.manual_seed(11)
    ref = oafno.AFNONet(**kw).eval()
    with torch.no_grad():
        for n, p in ref.named_parameters():      # away from the init scale: biases, LayerNorm affine and filter weights matter
            if "filter" in n:
                p.mul_(6.0)
            elif n.endswith("bias") or "norm" in n:
                p.add_(0.3 * torch.randn_like(p))
    const, presc, prog = synthetic(B, 4, 4, 1, Cg, H, W, seed=77)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
    outs = {}
    for prec in (native.PREC_FP32_TC, native.PREC_FP32):
        mod = pkg.FourCastNet(**kw).cuda().eval()
        mod.precision = prec
        mod.load_state_dict(ref.state_dict(), strict=True)
        with torch.no_grad():
            outs[prec] = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
        assert max_rel(outs[prec], want) < 2 * TOL_STEP
    assert max_rel(outs[native.PREC_FP32_TC], outs[native.PREC_FP32]) < 2 * TOL_STEP


@pytest.mark.parametrize("H,W,E", [(32, 64, 64), (8, 20, 48)])
def test_fourcastnet_fno2d_filter_matches_oracle(pkg, H, W, E):
    """``filter: FNO2D`` (fourcastnet.py:130-153): a one-layer neuralop FNO as the token mixer; the second case
    has a token grid smaller than the 12 x 7 stored modes."""
    from oracle import afno as oafno
    kw = dict(img_height=H, img_width=W, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
              prognostic_channels=3, embed_dim=E, depth=2, num_blocks=4, mlp_ratio=4.0, context_size=1,
              use_pos_embed=True, filter="FNO2D")
    torch.manual_seed(3)
    ref = oafno.AFNONet(**kw).eval()
    with torch.no_grad():
        for n, p in ref.named_parameters():
            if n.endswith("bias") or "norm" in n:
                p.add_(0.2 * torch.randn_like(p))
    mod = pkg.FourCastNet(**kw).cuda().eval()
    mod.load_state_dict(ref.state_dict(), strict=True)
    const, presc, prog = synthetic(3, 4, 4, 1, 3, H, W, seed=5)
    with torch.no_grad():
        want = ref(constants=const, prescribed=presc, prognostic=prog)
        got = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
    assert max_rel(got, want) < 2 * TOL_STEP


# ---------------------------------------------------------------------------------------
# host entry, metrics kernel, weight reload
# ---------------------------------------------------------------------------------------
@pytest.mark.parametrize("ctx,T,members", [(1, 4, None), (2, 7, None), (1, 4, 64), (2, 5, 66), (1, 2, 64), (1, 3, 256)])
def test_rollout_host_equals_device_rollout(pkg, ctx, T, members):
    # the host entry streams the later prescribed frames in under the earlier steps: same bits as the device entry
    # (64 members or more: the host entry runs the first and the last step in two parts of the members, 256 or more: in four)
    cls, kw, (B, H, W) = FNO_CASES["C2"]
    B = members or B
    kw = dict(kw, context_size=ctx)
    torch.manual_seed(1234)
    mod = getattr(pkg, cls)(**kw).cuda().eval()
    const, presc, prog = synthetic(B, T, 4, 1, 8, H, W)
    with torch.no_grad():
        dev = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
        host = mod.forward_host(constants=const.pin_memory(), prescribed=presc.pin_memory(),
                                prognostic=prog.pin_memory())
        again = mod.forward_host(constants=const.pin_memory(), prescribed=(presc + 1.0).pin_memory(),
                                 prognostic=prog.pin_memory())
    assert host.device.type == "cpu" and host.shape == (B, T - ctx, 8, H, W)
    assert torch.equal(dev, host)
    assert not torch.equal(again, host)      # a second call with other prescribed frames is not served stale ones


@pytest.mark.parametrize("members,no_forcings", [(64, False), (70, False), (64, True)])
def test_fourcastnet_rollout_host_equals_device_rollout(pkg, members, no_forcings):
    # the host entry's two half batches on the AFNO path (fused first / last kernels), with and without the
    # constants / prescribed groups (fourcastnet.py:294-307 skips absent groups)
    kw = dict(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=0 if no_forcings else 4,
              prescribed_channels=0 if no_forcings else 1, prognostic_channels=3, embed_dim=64, filter="AFNO2D", depth=2,
              mlp_ratio=4.0, num_blocks=4, sparsity_threshold=0.01, hard_thresholding_fraction=1.0, context_size=1,
              use_pos_embed=True)
    torch.manual_seed(7)
    mod = pkg.FourCastNet(**kw).cuda().eval()
    const, presc, prog = synthetic(members, 4, 4, 1, 3, 32, 64)
    if no_forcings:
        const = presc = None
    pin = lambda x: None if x is None else x.pin_memory()
    with torch.no_grad():
        dev = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).cpu()
        host = mod.forward_host(constants=pin(const), prescribed=pin(presc), prognostic=pin(prog))
    assert host.shape == (members, 3, 3, 32, 64)
    assert torch.equal(dev, host)


@pytest.mark.parametrize("B,T,Cn,H,W", [(5, 3, 4, 32, 64), (70, 2, 3, 32, 64), (3, 2, 2, 9, 15), (2, 1, 1, 181, 360)])
def test_metrics_kernel_matches_oracle(pkg, B, T, Cn, H, W):
    # float4 form (W % 4 == 0), more samples than one pass of a CTA covers, ragged planes (scalar form), sliced planes
    from oracle import metrics as om
    torch.manual_seed(2)
    o, t, c = torch.randn(B, T, Cn, H, W), torch.randn(B, T, Cn, H, W), 0.3 * torch.randn(Cn, H, W)
    rmse, acc = pkg.rollout_metrics(o.cuda(), t.cuda(), c.cuda())
    wr, wa = om.rmse_acc(o.double(), t.double(), c.double())
    assert max_rel(rmse, wr) < 1e-6
    assert float((acc.cpu() - wa).abs().max()) < 1e-6


@pytest.mark.parametrize("B,T,Cn,H,W", [(5, 3, 8, 32, 64), (2, 2, 3, 9, 15)])
def test_denormalize_split_matches_the_reference_loop(pkg, B, T, Cn, H, W):
    # evaluate.py:281-296 (x * std + mean per variable) + the per-variable arrays build_dataset slices out (:366-383)
    from dlwp_benchmark_b200 import evaluation as ev
    torch.manual_seed(4)
    x, mean, std = torch.randn(B, T, Cn, H, W), 10.0 * torch.randn(Cn), torch.rand(Cn) + 0.5
    got = ev.denormalize_split(x.cuda(), mean, std).cpu()
    want = (x * std.view(1, 1, Cn, 1, 1) + mean.view(1, 1, Cn, 1, 1)).permute(2, 0, 1, 3, 4)
    assert got.shape == (Cn, B, T, H, W)
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
    assert torch.equal(ev.denormalize_split(x.cuda()).cpu(), x.permute(2, 0, 1, 3, 4))


def test_weights_are_repacked_after_in_place_update(pkg):
    ref, mod, kw, (B, H, W) = _fno_pair(pkg, "C2")
    const, presc, prog = synthetic(B, 2, 4, 1, 8, H, W)
    with torch.no_grad():
        a = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog)).clone()
        for p in mod.parameters():
            p.mul_(1.5)
        for p in ref.parameters():
            p.mul_(1.5)
        b = mod(constants=_dev(const), prescribed=_dev(presc), prognostic=_dev(prog))
        want = ref(constants=const, prescribed=presc, prognostic=prog)
    assert not torch.equal(a, b)
    assert max_rel(b, want) < TOL_STEP
