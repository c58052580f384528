"""GPU: the training path of FNO2DModule / TFNO2DModule / FourCastNet (scripts/train.py:263-271 back-propagates through
the model): forward values and every gradient -- parameters (Tucker core and factors included) and inputs -- against autograd
through the CPU oracle evaluated in fp64, on one evaluation and on a multi-step rollout with context blending; then one
optimizer step of the reference's training-loop shape.

Stated tolerance: max|g - g_ref| / max|g_ref| <= 2e-4 per tensor (fp32 kernels, atomically accumulated weight gradients).
"""
import pytest
import torch

from helpers import max_rel, synthetic

pytestmark = pytest.mark.gpu

TOL_GRAD = 2e-4

CASES = {
    "fno-odd": ("FNO2DModule", dict(n_modes=[6, 10], constant_channels=1, prescribed_channels=1, prognostic_channels=2, hidden_channels=12,
                                    lifting_channels=24, projection_channels=20, n_layers=2, context_size=2), (3, 17, 30), 5),
    "tfno-c2": ("TFNO2DModule", dict(n_modes=[12, 12], constant_channels=4, prescribed_channels=1, prognostic_channels=8, hidden_channels=32,
                                     lifting_channels=256, projection_channels=256, n_layers=4, rank=0.8, context_size=1), (2, 32, 64), 3),
    # FourCastNet: the filter weights are scaled up (see _pair) so that ReLU and softshrink cut a real share of the entries
    "afno-c3": ("FourCastNet", dict(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
                                    prognostic_channels=8, embed_dim=64, depth=2, num_blocks=8, mlp_ratio=4.0, context_size=1,
                                    sparsity_threshold=0.1), (2, 32, 64), 3),
    "afno-patch": ("FourCastNet", dict(img_height=16, img_width=48, patch_size=[2, 4], constant_channels=2, prescribed_channels=0,
                                       prognostic_channels=3, embed_dim=48, depth=2, num_blocks=4, mlp_ratio=2.0, context_size=2,
                                       hard_thresholding_fraction=0.6, sparsity_threshold=0.05, use_pos_embed=False), (3, 16, 48), 4),
    "afno-fno2d": ("FourCastNet", dict(img_height=16, img_width=32, patch_size=[1, 1], constant_channels=1, prescribed_channels=1,
                                       prognostic_channels=2, filter="FNO2D", embed_dim=32, depth=2, mlp_ratio=2.0, context_size=1),
                   (2, 16, 32), 3),
}


@pytest.fixture(scope="module")
def pkg(built_lib):
    import dlwp_benchmark_b200 as p
    assert torch.cuda.is_available()
    return p


def _pair(pkg, case):
    from oracle import wrappers as owrap
    cls, kw, shape, T = CASES[case]
    torch.manual_seed(1234)
    if cls == "FourCastNet":
        from oracle import afno as oafno
        ref = oafno.AFNONet(**kw)
        with torch.no_grad():
            for name, p in ref.named_parameters():
                if ".filter.w" in name or ".filter.b" in name:
                    p.mul_(8.0)
                elif name.endswith("norm1.bias") or name.endswith("norm2.bias") or name == "pos_embed":
                    p.add_(0.1 * torch.randn_like(p))
        ref = ref.double()
        if kw.get("filter") == "FNO2D":      # oracle FNO2D.forward casts to fp32 as the reference does (:149); the fp64 checker skips the cast
            for blk in ref.blocks:
                blk.filter.forward = (lambda f: lambda x: f.fno(x.permute(0, 3, 1, 2)).permute(0, 2, 3, 1) + x)(blk.filter)
    else:
        ref = getattr(owrap, cls)(**kw).double()
    mod = getattr(pkg, cls)(**kw).cuda()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}, strict=True)
    return ref, mod, kw, shape, T


@pytest.mark.parametrize("case", list(CASES))
def test_rollout_gradients_match_oracle_autograd(pkg, case):
    ref, mod, kw, (B, H, W), T = _pair(pkg, case)
    const, presc, prog = synthetic(B, T, kw["constant_channels"], kw["prescribed_channels"], kw["prognostic_channels"], H, W, seed=3)
    ref_grads = dict(ref.named_parameters())
    tgt = synthetic(B, T - kw["context_size"], 0, 0, kw["prognostic_channels"], H, W, seed=9)[2]
    # reference: fp64 autograd through the oracle
    prog_r = prog.double().requires_grad_(True)
    out_r = ref(constants=None if const is None else const.double(), prescribed=None if presc is None else presc.double(), prognostic=prog_r)
    loss_r = (out_r - tgt.double()).pow(2).mean()
    loss_r.backward()
    # this library
    prog_g = prog.cuda().requires_grad_(True)
    out_g = mod(constants=None if const is None else const.cuda(), prescribed=None if presc is None else presc.cuda(), prognostic=prog_g)
    assert out_g.requires_grad and out_g.shape == out_r.shape
    assert max_rel(out_g.detach(), out_r.detach()) < 1e-5
    loss_g = (out_g - tgt.cuda()).pow(2).mean()
    loss_g.backward()
    assert abs(float(loss_g) - float(loss_r)) < 1e-5 * abs(float(loss_r))
    worst = ("", 0.0)
    for name, p in mod.named_parameters():
        if name.startswith("norm."):      # AFNONet.norm is held for the state_dict only (fourcastnet.py:263), no gradient on either side
            assert p.grad is None and ref_grads[name].grad is None
            continue
        g_ref = ref_grads[name].grad
        assert p.grad is not None, name
        err = max_rel(p.grad, g_ref)
        worst = max(worst, (name, err), key=lambda t: t[1])
        assert err < TOL_GRAD, (name, err)
    err_in = max_rel(prog_g.grad, prog_r.grad)
    print(f"{case}: worst parameter-gradient error {worst[1]:.2e} ({worst[0]}), input-gradient error {err_in:.2e}")
    assert err_in < TOL_GRAD


def test_one_training_step_of_the_reference_loop_shape(pkg):
    # scripts/train.py:263-271: output = model(...); loss = criterion(output, target); loss.backward(); optimizer.step()
    _, mod, kw, (B, H, W), T = _pair(pkg, "tfno-c2")
    const, presc, prog = (t.cuda() for t in synthetic(B, T, 4, 1, 8, H, W, seed=5))
    tgt = prog[:, 1:] * 0.9
    opt = torch.optim.Adam(mod.parameters(), lr=1e-3)
    losses = []
    for _ in range(4):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(mod(constants=const, prescribed=presc, prognostic=prog), tgt)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses))) and losses[-1] < losses[0]
    # the inference plan sees the updated weights afterwards (weights are re-uploaded when a parameter changed)
    with torch.no_grad():
        out = mod(constants=const, prescribed=presc, prognostic=prog)
        assert abs(float(torch.nn.functional.mse_loss(out, tgt)) - losses[-1]) < abs(losses[-1])


@pytest.mark.parametrize("C,nb,h,w,frac", [(192, 4, 12, 20, 1.0), (768, 16, 6, 10, 0.5), (40, 5, 9, 14, 1.0)])
def test_layernorm_and_afno_filter_ops_at_other_widths(pkg, C, nb, h, w, frac):
    """The two FourCastNet primitives alone: register-array widths of the LayerNorm backward kernel (C / 32 up to 24) and the
    1 / 4 / 16 entries-per-thread forms of the block weight-gradient kernel, against fp64 autograd through torch / the oracle."""
    from oracle import afno as oafno
    from dlwp_benchmark_b200 import training
    from dlwp_benchmark_b200.plan import afno_kept_region
    torch.manual_seed(C)
    B, lam = 2, 0.05
    x = torch.randn(B * h * w, C)
    gamma, beta = 1.0 + 0.2 * torch.randn(C), 0.2 * torch.randn(C)
    gy = torch.randn(B * h * w, C)
    # LayerNorm
    ref_in = [t.double().requires_grad_(True) for t in (x, gamma, beta)]
    torch.nn.functional.layer_norm(ref_in[0], (C,), ref_in[1], ref_in[2], 1e-6).backward(gy.double())
    got_in = [t.cuda().requires_grad_(True) for t in (x, gamma, beta)]
    y = training._LayerNorm.apply(got_in[0], got_in[1], got_in[2], 1e-6)
    y.backward(gy.cuda())
    for a, b, name in zip(got_in, ref_in, ("dx", "dgamma", "dbeta")):
        assert max_rel(a.grad, b.grad) < 2e-5, (name, max_rel(a.grad, b.grad))
    # AFNO2D
    filt = oafno.AFNO2D(C, nb, lam, frac)
    with torch.no_grad():
        for p in filt.parameters():
            p.mul_(4.0)
    filt = filt.double()
    xr = x.double().reshape(B, h, w, C).requires_grad_(True)
    er = gy.double().reshape(B, h, w, C).requires_grad_(True)
    out_r = filt(xr) + er
    out_r.backward(torch.cos(out_r.detach()))
    rows, cols = afno_kept_region(h, w // 2 + 1, frac)
    geom = (B, h, w, tuple(rows), cols, nb, lam)
    xg, eg = x.cuda().requires_grad_(True), gy.cuda().requires_grad_(True)
    par = [p.detach().float().cuda().requires_grad_(True) for p in (filt.w1, filt.b1, filt.w2, filt.b2)]
    out_g = training._AfnoFilter.apply(xg, eg, *par, geom)
    assert max_rel(out_g.detach().reshape(B, h, w, C), out_r.detach()) < 1e-5
    out_g.backward(torch.cos(out_r.detach()).float().cuda().reshape(B * h * w, C))
    assert max_rel(xg.grad.reshape(B, h, w, C), xr.grad) < TOL_GRAD
    assert torch.equal(eg.grad.cpu(), torch.cos(out_r.detach()).float().reshape(B * h * w, C))
    for a, b, name in zip(par, (filt.w1, filt.b1, filt.w2, filt.b2), ("w1", "b1", "w2", "b2")):
        assert max_rel(a.grad, b.grad) < TOL_GRAD, (name, max_rel(a.grad, b.grad))


def test_fourcastnet_training_steps(pkg):
    _, mod, kw, (B, H, W), T = _pair(pkg, "afno-c3")
    const, presc, prog = (t.cuda() for t in synthetic(B, T, 4, 1, 8, H, W, seed=5))
    tgt = prog[:, 1:] * 0.9
    opt = torch.optim.Adam(mod.parameters(), lr=1e-3)
    losses = []
    for _ in range(4):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(mod(constants=const, prescribed=presc, prognostic=prog), tgt)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses))) and losses[-1] < losses[0]
    with torch.no_grad():      # the tensor-core inference plan on the updated weights
        out = mod(constants=const, prescribed=presc, prognostic=prog)
        assert abs(float(torch.nn.functional.mse_loss(out, tgt)) - losses[-1]) < abs(losses[-1])


def test_device_prefetcher_feeds_the_training_loop(pkg):
    """the loader's (constants, prescribed, prognostic, target) batches arrive on the device ahead of the consumer,
    in order, bit-identical; a NaN tensor stands for an absent input (train.py:180-181) and arrives as None; a loader
    exception reaches the consumer; stopping early does not hang."""
    g = torch.Generator().manual_seed(0)
    batches = []
    for n in range(7):
        B = 2 + n % 3
        batches.append((torch.full((B, 1, 4, 8, 16), float("nan")) if n % 2 else torch.randn(B, 1, 4, 8, 16, generator=g),
                        torch.randn(B, 3, 1, 8, 16, generator=g), torch.randn(B, 3, 8, 8, 16, generator=g),
                        torch.randn(B, 2, 8, 8, 16, generator=g).double()))
    got = list(pkg.DevicePrefetcher(batches, "cuda", depth=2))
    assert len(got) == len(batches)
    for n, (want, have) in enumerate(zip(batches, got)):
        assert (have[0] is None) == bool(n % 2)
        for w, h in zip(want[1:], have[1:]):
            assert h.device.type == "cuda" and h.dtype == w.dtype and torch.equal(h.cpu(), w)
        if have[0] is not None:
            assert torch.equal(have[0].cpu(), want[0])
    # early stop
    it = iter(pkg.DevicePrefetcher(batches, "cuda", depth=1))
    next(it)
    it.close()

    def bad():
        yield batches[0]
        raise RuntimeError("loader failed")
    with pytest.raises(RuntimeError, match="loader failed"):
        list(pkg.DevicePrefetcher(bad(), "cuda"))
    # and the loop shape of train.py:176-271 on top of it
    _, mod, kw, (B, H, W), T = _pair(pkg, "tfno-c2")
    data = [synthetic(B, T, 4, 1, 8, H, W, seed=s) + (synthetic(B, T - 1, 0, 0, 8, H, W, seed=50 + s)[2],) for s in range(3)]
    opt = torch.optim.Adam(mod.parameters(), lr=1e-3)
    for constants, prescribed, prognostic, target in pkg.DevicePrefetcher(data, "cuda"):
        opt.zero_grad()
        loss = torch.nn.functional.mse_loss(mod(constants=constants, prescribed=prescribed, prognostic=prognostic), target)
        loss.backward()
        opt.step()
    assert torch.isfinite(loss)
