"""Drop-in ``SpectralConv2d`` and ``FourierResidualBlock`` (``src/dlwpbench/models/unet/unet.py:19-69,780-836``): same
constructor arguments, parameter names and shapes as the reference (so its ``state_dict`` loads), forward pass in
``libdlwp_b200.so`` (``dlwp_spectral_conv2d`` / ``dlwp_fourier_residual_block``).  Forward only, CUDA only.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import native


def _check(module: nn.Module, x: torch.Tensor):
    if x.device.type != "cuda":
        raise native.NativeError("inputs must live on a CUDA device (no CPU fallback)")
    if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in module.parameters())):
        raise NotImplementedError("forward-only: call under torch.no_grad()")


def _corner_rows(H: int, m1: int):
    return list(range(m1)) + list(range(H - m1, H))


class SpectralConv2d(nn.Module):
    """``unet.py:19-69``: rfft2, two corners of retained modes, einsum("bixy,ioxy->boxy"), irfft2."""

    def __init__(self, in_channels: int, out_channels: int, modes1: int, modes2: int):
        super().__init__()
        self.in_channels, self.out_channels, self.modes1, self.modes2 = in_channels, out_channels, modes1, modes2
        self.scale = 1 / (in_channels * out_channels)
        self.weights1 = nn.Parameter(self.scale * torch.rand(in_channels, out_channels, modes1, modes2, 2))
        self.weights2 = nn.Parameter(self.scale * torch.rand(in_channels, out_channels, modes1, modes2, 2))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _check(self, x)
        B, Ci, H, W = x.shape
        x = x.to(torch.float32).contiguous()
        # one weight tensor whose rows are the two corners (rows [:m1], then rows [-m1:])
        weight = torch.cat([self.weights1.detach(), self.weights2.detach()], dim=2).to(torch.float32).contiguous()
        rows = native.int32_array(_corner_rows(H, self.modes1))
        y = torch.empty(B, self.out_channels, H, W, device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            native.check(native.lib().dlwp_spectral_conv2d(
                C.c_void_p(x.data_ptr()), C.c_void_p(weight.data_ptr()), None, C.c_void_p(y.data_ptr()), B, Ci,
                self.out_channels, H, W, 2 * self.modes1, self.modes2, rows, rows, native.FFT_NORM_BACKWARD, st),
                "dlwp_spectral_conv2d")
        return y


class FourierResidualBlock(nn.Module):
    """``unet.py:780-836`` with the default GELU activation; ``norm=True`` uses GroupNorm(n_groups)."""

    def __init__(self, in_channels: int, out_channels: int, modes1: int = 16, modes2: int = 16, activation=None,
                 norm: bool = False, n_groups: int = 1):
        super().__init__()
        if activation is not None and not isinstance(activation, nn.GELU):
            raise NotImplementedError("the CUDA block implements the reference's default activation (exact GELU)")
        self.modes1, self.modes2, self.n_groups = modes1, modes2, n_groups
        self.fourier1 = SpectralConv2d(in_channels, out_channels, modes1, modes2)
        self.conv1 = nn.Conv2d(in_channels, out_channels, kernel_size=1)
        self.fourier2 = SpectralConv2d(out_channels, out_channels, modes1, modes2)
        self.conv2 = nn.Conv2d(out_channels, out_channels, kernel_size=1)
        self.shortcut = nn.Conv2d(in_channels, out_channels, kernel_size=(1, 1)) if in_channels != out_channels \
            else nn.Identity()
        self.norm1 = nn.GroupNorm(n_groups, in_channels) if norm else nn.Identity()
        self.norm2 = nn.GroupNorm(n_groups, out_channels) if norm else nn.Identity()
        self._workspace = None
        self._keep = None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        _check(self, x)
        B, Ci, H, W = x.shape
        Co = self.conv1.out_channels
        x = x.to(torch.float32).contiguous()
        keep = []

        def ptr(t):
            if t is None:
                return None
            t = t.detach().to(device=x.device, dtype=torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        w = native.FrbWeights()
        w.f1_w1, w.f1_w2 = ptr(self.fourier1.weights1), ptr(self.fourier1.weights2)
        w.conv1_w, w.conv1_b = ptr(self.conv1.weight), ptr(self.conv1.bias)
        w.f2_w1, w.f2_w2 = ptr(self.fourier2.weights1), ptr(self.fourier2.weights2)
        w.conv2_w, w.conv2_b = ptr(self.conv2.weight), ptr(self.conv2.bias)
        if isinstance(self.shortcut, nn.Conv2d):
            w.short_w, w.short_b = ptr(self.shortcut.weight), ptr(self.shortcut.bias)
        eps = 1e-5
        if isinstance(self.norm1, nn.GroupNorm):
            w.norm1_w, w.norm1_b = ptr(self.norm1.weight), ptr(self.norm1.bias)
            w.norm2_w, w.norm2_b = ptr(self.norm2.weight), ptr(self.norm2.bias)
            eps = self.norm1.eps
        w.n_groups, w.eps = self.n_groups, eps
        lib = native.lib()
        nbytes = C.c_size_t()
        native.check(lib.dlwp_fourier_residual_block_workspace(B, Ci, Co, H, W, self.modes1, self.modes2, C.byref(nbytes)),
                     "dlwp_fourier_residual_block_workspace")
        ws = self._workspace
        if ws is None or ws.numel() < nbytes.value or ws.device != x.device:
            ws = self._workspace = torch.empty(nbytes.value, dtype=torch.uint8, device=x.device)
        y = torch.empty(B, Co, H, W, device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            native.check(lib.dlwp_fourier_residual_block(
                C.c_void_p(x.data_ptr()), C.byref(w), C.c_void_p(y.data_ptr()), B, Ci, Co, H, W, self.modes1, self.modes2,
                C.c_void_p(ws.data_ptr()), ws.numel(), st), "dlwp_fourier_residual_block")
        self._keep = keep      # fp32 copies made for this call (none for fp32 parameters) live until the next call
        return y
