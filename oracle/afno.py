"""Oracle: restatement of the in-repo FourCastNet (``AFNONet``) forward pass.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  **Parity pinned**: this file is
checked against the reference source executed unmodified (``oracle/ref_loader.py`` in the
build container) and against ``tests/golden/afno_*.npz`` generated from it.

Follows ``/root/reference/src/dlwpbench/models/fourcastnet/fourcastnet.py``:
``Mlp`` :41-57, ``AFNO2D`` :60-127, ``Block`` :156-193, ``AFNONet`` :214-361 (rollout loop
in the working form of :489-527), ``PatchEmbed`` :530-543.  The arithmetic is written with
complex tensors instead of the reference's eight real einsums, so agreement with the
reference is evidence rather than a tautology.  Parameter names match the reference.
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def afno_kept_region(h: int, w_half: int, frac: float):
    """Rows / columns of the rfft2 spectrum the block MLP touches (fourcastnet.py:93-96).

    ``total = h // 2 + 1`` and ``kept = int(total * frac)`` are derived from the *row*
    count but ``kept`` also truncates the column axis.  Python slicing clamps both.
    """
    total = h // 2 + 1
    kept = int(total * frac)
    rows = list(range(h))[max(total - kept, 0): total + kept]
    cols = list(range(w_half))[:kept]
    return rows, cols


class AFNO2D(nn.Module):
    def __init__(self, hidden_size, num_blocks=8, sparsity_threshold=0.01, hard_thresholding_fraction=1,
                 hidden_size_factor=1):
        super().__init__()
        if hidden_size % num_blocks:
            raise AssertionError(f"hidden_size {hidden_size} should be divisble by num_blocks {num_blocks}")
        self.hidden_size, self.num_blocks = hidden_size, num_blocks
        self.block_size = hidden_size // num_blocks
        self.sparsity_threshold = sparsity_threshold
        self.hard_thresholding_fraction = hard_thresholding_fraction
        self.hidden_size_factor = hidden_size_factor
        bs, f = self.block_size, hidden_size_factor
        self.w1 = nn.Parameter(0.02 * torch.randn(2, num_blocks, bs, bs * f))
        self.b1 = nn.Parameter(0.02 * torch.randn(2, num_blocks, bs * f))
        self.w2 = nn.Parameter(0.02 * torch.randn(2, num_blocks, bs * f, bs))
        self.b2 = nn.Parameter(0.02 * torch.randn(2, num_blocks, bs))

    def forward(self, x):
        resid = x
        in_dtype = x.dtype
        x = x.float() if x.dtype != torch.float64 else x       # reference forces fp32 (:81-82)
        B, H, W, C = x.shape
        wh = W // 2 + 1
        cdt = torch.complex64 if x.dtype == torch.float32 else torch.complex128
        spec = torch.fft.rfft2(x, dim=(1, 2), norm="ortho").reshape(B, H, wh, self.num_blocks, self.block_size)
        rows, cols = afno_kept_region(H, wh, self.hard_thresholding_fraction)
        out = torch.zeros(B, H, wh, self.num_blocks, self.block_size, dtype=cdt, device=x.device)
        if rows and cols:
            r0, r1, c1 = rows[0], rows[-1] + 1, cols[-1] + 1
            w1 = torch.complex(self.w1[0], self.w1[1]).to(cdt)
            w2 = torch.complex(self.w2[0], self.w2[1]).to(cdt)
            b1 = torch.complex(self.b1[0], self.b1[1]).to(cdt)
            b2 = torch.complex(self.b2[0], self.b2[1]).to(cdt)
            z = torch.einsum("...ki,kio->...ko", spec[:, r0:r1, :c1], w1) + b1
            z = torch.complex(F.relu(z.real), F.relu(z.imag))      # ReLU on re and im separately
            z = torch.einsum("...ki,kio->...ko", z, w2) + b2
            out[:, r0:r1, :c1] = z
        lam = self.sparsity_threshold
        out = torch.complex(F.softshrink(out.real, lam), F.softshrink(out.imag, lam))
        y = torch.fft.irfft2(out.reshape(B, H, wh, C), s=(H, W), dim=(1, 2), norm="ortho")
        return y.to(in_dtype) + resid


class FNO2D(nn.Module):
    """``fourcastnet.py:130-153``: a one-layer neuralop ``FNO`` (hidden = lifting = projection = embed_dim,
    ``n_modes=(12, 12)``) as the token mixer, selected by ``filter: FNO2D`` (``eval`` at ``:259``).  ``Block`` passes
    ``(dim, num_blocks, sparsity_threshold, hard_thresholding_fraction)`` positionally (``:173``); only ``dim`` is
    used.  The FNO itself is the restatement of ``oracle/neuralop_fno.py`` (parity unpinned, see there)."""

    def __init__(self, hidden_size, num_blocks=8, sparsity_threshold=0.01, hard_thresholding_fraction=1,
                 hidden_size_factor=1, n_modes=(12, 12), max_n_modes=None, n_layers=1):
        super().__init__()
        from .neuralop_fno import FNO
        self.fno = FNO(n_modes=n_modes, max_n_modes=max_n_modes, in_channels=hidden_size, hidden_channels=hidden_size,
                       lifting_channels=hidden_size, projection_channels=hidden_size, out_channels=hidden_size,
                       n_layers=n_layers)

    def forward(self, x):                                   # [B, h, w, C]
        bias = x
        y = self.fno(x.float().permute(0, 3, 1, 2)).permute(0, 2, 3, 1)
        return y + bias


class Mlp(nn.Module):
    def __init__(self, in_features, hidden_features):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.fc2 = nn.Linear(hidden_features, in_features)

    def forward(self, x):
        return self.fc2(F.gelu(self.fc1(x)))


class Block(nn.Module):
    def __init__(self, dim, mlp_ratio, num_blocks, sparsity_threshold, hard_thresholding_fraction, filter="AFNO2D"):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.filter = {"AFNO2D": AFNO2D, "FNO2D": FNO2D}[filter](dim, num_blocks, sparsity_threshold,
                                                                 hard_thresholding_fraction)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = Mlp(dim, int(dim * mlp_ratio))

    def forward(self, x):
        x = self.filter(self.norm1(x)) + x          # double_skip=True (default, :167,185-187)
        return self.mlp(self.norm2(x)) + x


class PatchEmbed(nn.Module):
    def __init__(self, img_size, patch_size, in_chans, embed_dim):
        super().__init__()
        self.img_size = tuple(img_size)
        self.patch_size = (int(patch_size[0]), int(patch_size[1]))
        self.num_patches = (img_size[1] // self.patch_size[1]) * (img_size[0] // self.patch_size[0])
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=self.patch_size, stride=self.patch_size)

    def forward(self, x):
        if tuple(x.shape[-2:]) != self.img_size:
            raise AssertionError(f"Input image size {tuple(x.shape[-2:])} doesn't match model {self.img_size}.")
        return self.proj(x).flatten(2).transpose(1, 2)


class AFNONet(nn.Module):
    """FourCastNet.  Constructor keywords = keys of ``configs/model/fourcastnet.yaml``."""

    def __init__(self, img_height=720, img_width=1440, patch_size=(16, 16), constant_channels=4,
                 prescribed_channels=0, prognostic_channels=1, filter="AFNO2D", embed_dim=768, depth=12,
                 mlp_ratio=4.0, drop_rate=0.0, drop_path_rate=0.0, num_blocks=16, sparsity_threshold=0.01,
                 hard_thresholding_fraction=1.0, context_size=1, use_pos_embed=True, **kwargs):
        super().__init__()
        if filter not in ("AFNO2D", "FNO2D"):
            raise NotImplementedError("oracle restates the AFNO2D and FNO2D filters")
        if drop_rate or drop_path_rate:
            raise NotImplementedError("forward-only oracle: dropout / drop-path must be 0")
        self.img_size = (img_height, img_width)
        self.patch_size = (int(patch_size[0]), int(patch_size[1]))
        self.in_chans = constant_channels + (prescribed_channels + prognostic_channels) * context_size
        self.out_chans = prognostic_channels
        self.embed_dim, self.context_size, self.use_pos_embed = embed_dim, context_size, use_pos_embed
        self.patch_embed = PatchEmbed(self.img_size, self.patch_size, self.in_chans, embed_dim)
        if use_pos_embed:
            self.pos_embed = nn.Parameter(torch.zeros(1, self.patch_embed.num_patches, embed_dim))
            nn.init.trunc_normal_(self.pos_embed, std=0.02)
        self.h = img_height // self.patch_size[0]
        self.w = img_width // self.patch_size[1]
        self.blocks = nn.ModuleList(
            Block(embed_dim, mlp_ratio, num_blocks, sparsity_threshold, hard_thresholding_fraction, filter)
            for _ in range(depth))
        self.norm = nn.LayerNorm(embed_dim, eps=1e-6)       # defined, never applied (:263 vs :283-293)
        self.head = nn.Linear(embed_dim, self.out_chans * self.patch_size[0] * self.patch_size[1], bias=False)
        for m in self.modules():                             # _init_weights (:270-277): nn.Linear and LayerNorm only
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)

    def backbone(self, x):
        """One model evaluation ``[B, Cin, H, W] -> [B, Cg, H, W]`` (:345-357)."""
        B = x.shape[0]
        t = self.patch_embed(x)
        if self.use_pos_embed:
            t = t + self.pos_embed
        t = t.reshape(B, self.h, self.w, self.embed_dim)
        for blk in self.blocks:
            t = blk(t)
        t = self.head(t)                                     # [B, h, w, p1*p2*Cg]
        p1, p2 = self.patch_size
        t = t.reshape(B, self.h, self.w, p1, p2, self.out_chans)
        return t.permute(0, 5, 1, 3, 2, 4).reshape(B, self.out_chans, self.h * p1, self.w * p2)

    def forward(self, constants=None, prescribed=None, prognostic=None):
        from .wrappers import rollout
        return rollout(self.backbone, self.context_size, constants, prescribed, prognostic)
