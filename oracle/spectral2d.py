"""Oracle: restatement of the in-repo dense Fourier layer and residual block.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  **Parity pinned** against the
reference source (``oracle/ref_loader.py``; ``tests/golden/spectral2d_*.npz``).

Follows ``/root/reference/src/dlwpbench/models/unet/unet.py``: ``batchmul2d`` :15-17,
``SpectralConv2d`` :19-69 (rfft2 with the default "backward" norm, two row corners
``[:m1]`` / ``[-m1:]``, columns ``[:m2]``), ``FourierResidualBlock`` :780-836.
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class SpectralConv2d(nn.Module):
    def __init__(self, in_channels, out_channels, modes1, modes2):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.modes1, self.modes2 = modes1, modes2
        scale = 1.0 / (in_channels * out_channels)
        self.weights1 = nn.Parameter(scale * torch.rand(in_channels, out_channels, modes1, modes2, 2))
        self.weights2 = nn.Parameter(scale * torch.rand(in_channels, out_channels, modes1, modes2, 2))

    def forward(self, x):
        B, _, H, W = x.shape
        m1, m2 = self.modes1, self.modes2
        xf = torch.fft.rfft2(x)
        out = torch.zeros(B, self.out_channels, H, W // 2 + 1, dtype=xf.dtype, device=x.device)
        lo = torch.view_as_complex(self.weights1.contiguous()).to(xf.dtype)
        hi = torch.view_as_complex(self.weights2.contiguous()).to(xf.dtype)
        out[:, :, :m1, :m2] = torch.einsum("bixy,ioxy->boxy", xf[:, :, :m1, :m2], lo)
        out[:, :, -m1:, :m2] = torch.einsum("bixy,ioxy->boxy", xf[:, :, -m1:, :m2], hi)
        return torch.fft.irfft2(out, s=(H, W))


class FourierResidualBlock(nn.Module):
    """act(norm1) -> (spectral + 1x1) -> act(norm2) -> (spectral + 1x1) + shortcut (``unet.py:780-836``); act = exact
    GELU (the constructor default), norm = GroupNorm(n_groups) or Identity."""

    def __init__(self, in_channels, out_channels, modes1=16, modes2=16, norm=False, n_groups=1):
        super().__init__()
        self.fourier1 = SpectralConv2d(in_channels, out_channels, modes1, modes2)
        self.conv1 = nn.Conv2d(in_channels, out_channels, 1)
        self.fourier2 = SpectralConv2d(out_channels, out_channels, modes1, modes2)
        self.conv2 = nn.Conv2d(out_channels, out_channels, 1)
        self.shortcut = nn.Conv2d(in_channels, out_channels, 1) if in_channels != out_channels else nn.Identity()
        self.norm1 = nn.GroupNorm(n_groups, in_channels) if norm else nn.Identity()
        self.norm2 = nn.GroupNorm(n_groups, out_channels) if norm else nn.Identity()

    def forward(self, x):
        h = F.gelu(self.norm1(x))
        out = F.gelu(self.norm2(self.fourier1(h) + self.conv1(h)))
        return self.fourier2(out) + self.conv2(out) + self.shortcut(x)
