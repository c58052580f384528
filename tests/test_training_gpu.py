"""GPU: the training path of FNO2DModule / TFNO2DModule (scripts/train.py:263-271 back-propagates through
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
    ref = getattr(owrap, cls)(**kw).double()
    mod = getattr(pkg, cls)(**kw).cuda()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}, strict=True)
    return ref, mod, kw, shape, T


@pytest.mark.parametrize("case", list(CASES))
def test_rollout_gradients_match_oracle_autograd(pkg, case):
    ref, mod, kw, (B, H, W), T = _pair(pkg, case)
    const, presc, prog = synthetic(B, T, kw["constant_channels"], kw["prescribed_channels"], kw["prognostic_channels"], H, W, seed=3)
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
    ref_grads = dict(ref.named_parameters())
    for name, p in mod.named_parameters():
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
