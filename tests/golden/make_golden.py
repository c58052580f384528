"""Generate the committed golden fixtures by running the REAL reference source.

Run in the build container only (needs ``/root/reference``):
    python tests/golden/make_golden.py
Writes ``tests/golden/*.npz``.  Each file holds the constructor kwargs (JSON), the
reference module's ``state_dict`` (keys prefixed ``sd/``), seeded inputs and the outputs the
unmodified reference code produced for them on CPU in fp32.

AFNONet: outputs come from ``AFNONet.forward`` itself for one step (the fork's own loop
crashes beyond one step, fourcastnet.py:336-340) and from ``forward_features`` + ``head`` +
rearrange driven step by step with the working upstream loop (fourcastnet.py:502-525) for
multi-step fixtures.  SpectralConv2d / FourierResidualBlock: ``unet.py:14-69,780-836``.
FNO / TFNO have no fixture here: the reference's arithmetic for them is in the absent
``neuralop`` package (parity unpinned, see oracle/neuralop_fno.py).
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import ref_loader  # noqa: E402


def _sd(module):
    return {"sd/" + k: v.detach().numpy() for k, v in module.state_dict().items()}


def _randomise(model, gen):
    """The reference inits biases to 0 and LN to (1, 0); perturb so every term is exercised."""
    with torch.no_grad():
        for n, p in model.named_parameters():
            if n.endswith("bias") or "norm" in n:
                p.add_(0.1 * torch.randn(p.shape, generator=gen))
            if "filter" in n:
                p.mul_(6.0)       # lift the spectral MLP above the softshrink threshold


def afno_case(name, kw, B, steps, seed):
    ref = ref_loader.load_fourcastnet()
    torch.manual_seed(seed)
    gen = torch.Generator().manual_seed(seed)
    model = ref.AFNONet(**kw).eval()
    _randomise(model, gen)
    ctx = kw["context_size"]
    H, W = kw["img_height"], kw["img_width"]
    T = ctx + steps
    const = torch.randn(B, 1, kw["constant_channels"], H, W, generator=gen)
    presc = torch.randn(B, T, kw["prescribed_channels"], H, W, generator=gen)
    prog = torch.randn(B, T, kw["prognostic_channels"], H, W, generator=gen)
    with torch.no_grad():
        one = model(constants=const, prescribed=presc[:, :ctx + 1], prognostic=prog[:, :ctx + 1])
        outs = []
        for t in range(ctx, T):
            if t == ctx:
                pt = prog[:, 0:ctx]
            else:
                pt = torch.cat([prog[:, t - ctx:ctx], torch.stack(outs, 1)[:, -ctx:]], 1)
            x = model._prepare_inputs(constants=const, prescribed=presc[:, t - ctx:t], prognostic=pt)
            outs.append(pt[:, -1] + ref_loader.reference_afno_backbone(model, x))
        roll = torch.stack(outs, 1)
    assert torch.equal(one[:, 0], roll[:, 0])
    np.savez_compressed(os.path.join(HERE, name + ".npz"), kwargs=json.dumps(kw), constants=const.numpy(),
                        prescribed=presc.numpy(), prognostic=prog.numpy(), one_step=one.numpy(),
                        rollout=roll.numpy(), **_sd(model))
    print(name, "rollout", tuple(roll.shape), "absmax", float(roll.abs().max()))


def spectral_case(seed=7):
    u = ref_loader.load_unet_spectral()
    torch.manual_seed(seed)
    conv = u.SpectralConv2d(8, 12, 6, 5)
    blk = u.FourierResidualBlock(8, 12, modes1=6, modes2=5)
    x = torch.randn(2, 8, 16, 32)
    with torch.no_grad():
        y_conv, y_blk = conv(x), blk(x)
    out = {"x": x.numpy(), "y_conv": y_conv.numpy(), "y_block": y_blk.numpy()}
    out.update({"conv/" + k: v.detach().numpy() for k, v in conv.state_dict().items()})
    out.update({"block/" + k: v.detach().numpy() for k, v in blk.state_dict().items()})
    np.savez_compressed(os.path.join(HERE, "spectral2d.npz"), **out)
    print("spectral2d", tuple(y_conv.shape), tuple(y_blk.shape))
    # a second block with GroupNorm (norm=True, two groups) and the identity shortcut (in_channels == out_channels);
    # its own fixture file, drawn after the first one so that spectral2d.npz keeps its bytes
    nblk = u.FourierResidualBlock(12, 12, modes1=5, modes2=7, norm=True, n_groups=2)
    with torch.no_grad():
        for n_, p_ in nblk.named_parameters():
            if n_.startswith("norm"):
                p_.add_(0.3 * torch.randn_like(p_))
    xn = torch.randn(3, 12, 24, 40)
    with torch.no_grad():
        yn = nblk(xn)
    outn = {"x": xn.numpy(), "y_block": yn.numpy()}
    outn.update({"block/" + k: v.detach().numpy() for k, v in nblk.state_dict().items()})
    np.savez_compressed(os.path.join(HERE, "spectral2d_norm.npz"), **outn)
    print("spectral2d_norm", tuple(yn.shape))


if __name__ == "__main__":
    assert ref_loader.available(), "needs /root/reference"
    base = dict(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
                prognostic_channels=3, embed_dim=64, filter="AFNO2D", depth=4, mlp_ratio=4.0, drop_rate=0.0,
                drop_path_rate=0.0, num_blocks=4, sparsity_threshold=0.01, hard_thresholding_fraction=1.0,
                context_size=1, use_pos_embed=True)
    # configs/model/fourcastnet.yaml at 32x64, 1 sample, 3 steps
    afno_case("afno_c3", base, B=1, steps=3, seed=1234)
    # small: context 2, hard thresholding < 1, no pos-embed
    afno_case("afno_small", dict(base, img_height=16, img_width=32, embed_dim=32, depth=2, num_blocks=4,
                                 hard_thresholding_fraction=0.5, context_size=2, use_pos_embed=False,
                                 prognostic_channels=2), B=2, steps=4, seed=11)
    # patches 2x4
    afno_case("afno_patch", dict(base, img_height=16, img_width=32, patch_size=[2, 4], embed_dim=32, depth=2,
                                 num_blocks=8, mlp_ratio=2.0), B=2, steps=2, seed=12)
    spectral_case()
