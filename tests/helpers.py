"""Shared helpers for the parity tests (oracle = checker, never the thing under test)."""
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    out = {k: z[k] for k in z.files}
    if "kwargs" in out:
        out["kwargs"] = json.loads(str(out["kwargs"]))
    return out


def golden_state_dict(g, prefix="sd/"):
    return {k[len(prefix):]: torch.from_numpy(v.copy()) for k, v in g.items() if k.startswith(prefix)}


def max_rel(a, b):
    """max|a-b| / max|b|  -- the tolerance form BASELINE.json's north_star states."""
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def synthetic(B, T, Cc, Cp, Cg, H, W, seed=1234, smooth=False):
    g = torch.Generator().manual_seed(seed)

    def field(*shape):
        x = torch.randn(*shape, generator=g)
        if smooth:      # band-limit to |k| <= 8 and rescale to unit variance
            f = torch.fft.rfft2(x)
            ky = torch.fft.fftfreq(shape[-2], 1.0 / shape[-2]).abs().view(-1, 1)
            kx = torch.arange(shape[-1] // 2 + 1).view(1, -1)
            f = f * ((ky <= 8) & (kx <= 8))
            x = torch.fft.irfft2(f, s=shape[-2:])
            x = x / x.std()
        return x

    const = field(B, 1, Cc, H, W) if Cc else None
    presc = field(B, T, Cp, H, W) if Cp else None
    prog = field(B, T, Cg, H, W)
    return const, presc, prog
