"""Drop-in ``nn.Module`` front ends: same class names, constructor keywords, ``forward``
signature and ``state_dict`` keys as the reference, compute in ``libdlwp_b200.so``.

Reference: ``src/dlwpbench/models/fno/fno.py:12-146`` (FNO2DModule / TFNO2DModule wrapping
neuralop's FNO / TFNO) and ``src/dlwpbench/models/fourcastnet/fourcastnet.py:214-361``
(AFNONet, exported as ``FourCastNet``).  The modules below only HOLD parameters (with the
reference's names and shapes); every forward pass is one ``dlwp_rollout`` call.

Deliberate deviations: the autoregressive loop uses the working upstream form
(fourcastnet.py:502-506) instead of the fork's crashing ``[..].to()`` (fno.py:91-95); outputs
stay on the input device (the fork calls ``.cpu()`` per step, fno.py:104); forward-only.
"""
from __future__ import annotations

import math
from typing import Iterable, List, Optional, Tuple

import torch
import torch.nn as nn

from . import native
from .plan import NativePlan, afno_kept_region, centre_crop, neuralop_row_bins


# ======================================================================================
# shared forward machinery
# ======================================================================================
class _NativeRollout(nn.Module):
    """Base: lazily builds a native plan per (device, H, W), re-uploads weights when any
    parameter changed, and runs the rollout on torch's current stream."""

    # split-bf16 tcgen05 kernels where the shape is covered (fp32-parity: <= 1e-4), fp32 FMA kernels elsewhere
    precision = native.PREC_FP32_TC

    def _init_native(self):
        self._plans = {}

    # -- to be provided by subclasses ---------------------------------------------------
    def _desc(self, H: int, W: int):  # -> (fields, rows_in, rows_out)
        raise NotImplementedError

    def _native_weights(self, H: int, W: int) -> Iterable[Tuple[str, torch.Tensor]]:
        raise NotImplementedError

    # -- plumbing -------------------------------------------------------------------------
    def _weights_key(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def _plan_for(self, device, B, H, W) -> NativePlan:
        key = (device.index if device.index is not None else torch.cuda.current_device(), H, W, self.precision)
        plan = self._plans.get(key)
        if plan is None or plan.max_batch < B:
            if plan is not None:
                plan.close()
            fields, rows_in, rows_out = self._desc(H, W)
            fields = dict(fields, max_batch=B, height=H, width=W, precision=self.precision)
            plan = NativePlan(fields, rows_in, rows_out, device)
            self._plans[key] = plan
        wk = self._weights_key()
        if plan.weights_key != wk:
            plan.load(self._native_weights(H, W))
            plan.weights_key = wk
        return plan

    def _check(self, constants, prescribed, prognostic):
        if prognostic is None or prognostic.dim() != 5:
            raise ValueError("prognostic must be a [B, T, C, H, W] tensor")
        if self._needs_grad(constants, prescribed, prognostic) and not self._trainable:
            raise NotImplementedError(
                "this model family has a forward (rollout) pass only in dlwp_benchmark_b200; call it under "
                "torch.no_grad() / torch.inference_mode() as scripts/evaluate.py:205 does")
        B, T, Cg, H, W = prognostic.shape
        if T <= self.context_size:
            raise ValueError(f"sequence length {T} must exceed context_size {self.context_size}")
        if Cg != self.prognostic_channels:
            raise ValueError(f"prognostic has {Cg} channels, model expects {self.prognostic_channels}")
        for name, t, ch, frames in (("constants", constants, self.constant_channels, 1),
                                    ("prescribed", prescribed, self.prescribed_channels, T)):
            if t is None:
                if ch:
                    raise ValueError(f"{name} is None but the model was built with {ch} {name} channels")
                continue
            if tuple(t.shape) != (B, frames, ch, H, W):
                raise ValueError(f"{name} has shape {tuple(t.shape)}, expected {(B, frames, ch, H, W)}")
        return B, T, Cg, H, W

    # training (scripts/train.py:263-271): the modules route to training.py under grad mode
    _trainable = False

    def _needs_grad(self, *tensors):
        return torch.is_grad_enabled() and (any(p.requires_grad for p in self.parameters()) or any(
            t is not None and t.requires_grad for t in tensors))

    @staticmethod
    def _f32c(t):
        return None if t is None else t.to(torch.float32).contiguous()

    def forward(self, constants: torch.Tensor = None, prescribed: torch.Tensor = None,
                prognostic: torch.Tensor = None) -> torch.Tensor:
        B, T, Cg, H, W = self._check(constants, prescribed, prognostic)
        if prognostic.device.type != "cuda":
            raise native.NativeError("inputs must live on a CUDA device (no CPU fallback); "
                                     "use forward_host() for host buffers")
        dev = prognostic.device
        if self._needs_grad(constants, prescribed, prognostic):      # autograd path on the fp32 training primitives
            from . import training
            with torch.cuda.device(dev):
                return training.rollout(self, constants, prescribed, prognostic)
        with torch.cuda.device(dev):
            plan = self._plan_for(dev, B, H, W)
            c, p, g = self._f32c(constants), self._f32c(prescribed), self._f32c(prognostic)
            outs = torch.empty(B, T - self.context_size, Cg, H, W, device=dev, dtype=torch.float32)
            plan.rollout(c, p, g, outs)
        return outs

    def forward_host(self, constants=None, prescribed=None, prognostic=None, out=None, device=None):
        """Same rollout for HOST tensors (ideally pinned): H2D copies, rollout and the D2H copy
        of the result all happen inside ``dlwp_rollout_host``; returns a host tensor."""
        B, T, Cg, H, W = self._check(constants, prescribed, prognostic)
        if prognostic.device.type != "cpu":
            raise ValueError("forward_host takes host tensors")
        if self._needs_grad(constants, prescribed, prognostic):
            raise NotImplementedError("forward_host is the inference entry: call it under torch.no_grad()")
        dev = torch.device(device) if device is not None else next(self.parameters()).device
        if dev.type != "cuda":
            raise native.NativeError("move the module to a CUDA device first (no CPU fallback)")
        with torch.cuda.device(dev):
            plan = self._plan_for(dev, B, H, W)
            c, p, g = self._f32c(constants), self._f32c(prescribed), self._f32c(prognostic)
            if out is None:
                out = torch.empty(B, T - self.context_size, Cg, H, W, dtype=torch.float32,
                                  pin_memory=True)
            plan.rollout_host(c, p, g, out)
        return out

    def backbone(self, x: torch.Tensor) -> torch.Tensor:
        """One model evaluation on a prepared input ``[B, Cin, H, W]`` (``self.fno(x_t)``)."""
        if self._needs_grad(x):
            if not self._trainable:
                raise NotImplementedError("forward-only: call under torch.no_grad()")
            from . import training
            with torch.cuda.device(x.device):
                return training.backbone(self, x)
        B, _, H, W = x.shape
        with torch.cuda.device(x.device):
            plan = self._plan_for(x.device, B, H, W)
            y = torch.empty(B, self.prognostic_channels, H, W, device=x.device, dtype=torch.float32)
            plan.step(self._f32c(x), y)
        return y


# ======================================================================================
# FNO / TFNO parameter containers (neuralop 0.3.x names; they never run a forward pass)
# ======================================================================================
class _PointwiseMLP(nn.Module):
    """neuralop ``MLP``: ``fcs`` = 1x1 convolutions, GELU between them."""

    def __init__(self, in_channels, out_channels, hidden_channels, n_layers=2):
        super().__init__()
        widths = [in_channels] + [hidden_channels] * (n_layers - 1) + [out_channels]
        self.fcs = nn.ModuleList(nn.Conv2d(widths[i], widths[i + 1], 1) for i in range(n_layers))


class _DenseWeight(nn.Module):
    """Complex ``[Ci, Co, m1, m2']`` as a real view (tltorch dense tensor, parameter ``tensor``)."""

    def __init__(self, shape, std):
        super().__init__()
        self.tensor = nn.Parameter(torch.randn(*shape, 2) * (std / math.sqrt(2.0)))


def _tucker_ranks(shape, rank) -> List[int]:
    """Per-mode ranks so that core + factors hold about ``rank * prod(shape)`` values."""
    if isinstance(rank, (list, tuple)):
        return [int(r) for r in rank]
    if isinstance(rank, int) and not isinstance(rank, bool):
        return [min(rank, s) for s in shape]
    shape = [int(s) for s in shape]
    dense, squares, order = float(math.prod(shape)), float(sum(s * s for s in shape)), len(shape)
    lo, hi = 0.0, max(float(rank), 1.0)
    for _ in range(200):
        mid = 0.5 * (lo + hi)
        if dense * mid ** order + squares * mid - rank * dense > 0:
            hi = mid
        else:
            lo = mid
    f = 0.5 * (lo + hi)
    return [max(int(round(s * f)), 1) for s in shape]


class _Factors(nn.Module):
    def __init__(self, factors):
        super().__init__()
        for i, f in enumerate(factors):
            self.register_parameter(f"factor_{i}", nn.Parameter(f))
        self.n = len(factors)

    def __getitem__(self, i):
        return getattr(self, f"factor_{i}")


class _TuckerWeight(nn.Module):
    """Complex Tucker weight (tltorch): ``core`` + ``factors.factor_k``, all as real views."""

    def __init__(self, shape, rank, std):
        super().__init__()
        ranks = _tucker_ranks(shape, rank)
        self.rank = tuple(ranks)
        r = math.prod(math.sqrt(v) for v in ranks)
        std_f = (std / r) ** (1.0 / (len(shape) + 1)) / math.sqrt(2.0)
        self.core = nn.Parameter(torch.randn(*ranks, 2) * std_f)
        self.factors = _Factors([torch.randn(s, q, 2) * std_f for s, q in zip(shape, ranks)])


class _SpectralConvs(nn.Module):
    def __init__(self, channels, modes, n_layers, factorization, rank):
        super().__init__()
        std = (2.0 / (channels + channels)) ** 0.5
        shape = (channels, channels, *modes)
        if factorization is None:
            self.weight = nn.ModuleList(_DenseWeight(shape, std) for _ in range(n_layers))
        else:
            self.weight = nn.ModuleList(_TuckerWeight(shape, rank, std) for _ in range(n_layers))
        self.bias = nn.Parameter(std * torch.randn(n_layers, channels, 1, 1))


class _FNOBlocks(nn.Module):
    def __init__(self, channels, modes, n_layers, factorization, rank):
        super().__init__()
        self.convs = _SpectralConvs(channels, modes, n_layers, factorization, rank)
        self.fno_skips = nn.ModuleList(nn.Conv2d(channels, channels, 1, bias=False) for _ in range(n_layers))


class _FNOParams(nn.Module):
    def __init__(self, in_channels, hidden_channels, out_channels, lifting_channels, projection_channels,
                 n_layers, stored_modes, factorization, rank):
        super().__init__()
        self.fno_blocks = _FNOBlocks(hidden_channels, stored_modes, n_layers, factorization, rank)
        if lifting_channels:
            self.lifting = _PointwiseMLP(in_channels, hidden_channels, lifting_channels, 2)
        else:
            self.lifting = _PointwiseMLP(in_channels, hidden_channels, hidden_channels, 1)
        self.projection = _PointwiseMLP(hidden_channels, out_channels, projection_channels, 2)

    def forward(self, *a, **k):
        raise RuntimeError("parameter container only; call the wrapping module")


class FNO2DModule(_NativeRollout):
    """``models/fno/fno.py:12-106``.  Keyword surface identical to the reference; extra yaml
    keys (``type``, ``name``) are absorbed by ``**kwargs`` exactly as there."""

    _factorization = None
    _trainable = True

    def __init__(self, n_modes: list = [12, 12], constant_channels: int = 4, prescribed_channels: int = 1,
                 prognostic_channels: int = 8, hidden_channels: int = 32, lifting_channels: int = 256,
                 projection_channels: int = 256, n_layers: int = 4, max_n_modes: int = None, bias: bool = True,
                 context_size: int = 10, **kwargs):
        super().__init__()
        self._init_native()
        self._configure(n_modes, constant_channels, prescribed_channels, prognostic_channels, hidden_channels,
                        lifting_channels, projection_channels, n_layers, max_n_modes, context_size, rank=1.0)

    def _configure(self, n_modes, constant_channels, prescribed_channels, prognostic_channels, hidden_channels,
                   lifting_channels, projection_channels, n_layers, max_n_modes, context_size, rank):
        self.context_size = int(context_size)
        self.constant_channels = int(constant_channels)
        self.prescribed_channels = int(prescribed_channels)
        self.prognostic_channels = int(prognostic_channels)
        self.hidden_channels, self.n_layers = int(hidden_channels), int(n_layers)
        self.lifting_channels = int(lifting_channels) if lifting_channels else 0
        self.projection_channels = int(projection_channels)
        self.n_modes = [int(n_modes[0]), int(n_modes[1])]
        mx = self.n_modes if max_n_modes is None else [int(max_n_modes[0]), int(max_n_modes[1])]
        self.max_n_modes = mx
        self._stored = (mx[0], mx[1] // 2 + 1)             # rows, half-spectrum columns held by the weights
        in_channels = constant_channels + (prescribed_channels + prognostic_channels) * context_size
        self.in_channels = int(in_channels)
        self.fno = _FNOParams(self.in_channels, self.hidden_channels, self.prognostic_channels,
                              self.lifting_channels, self.projection_channels, self.n_layers, self._stored,
                              self._factorization, rank)

    # -- native description -------------------------------------------------------------
    def _modes_for(self, H, W):
        m1 = min(self.n_modes[0], self._stored[0], H)
        m2 = min(self.n_modes[1] // 2 + 1, self._stored[1], W // 2 + 1)
        return m1, m2

    def _desc(self, H, W):
        m1, m2 = self._modes_for(H, W)
        rows_in, rows_out = neuralop_row_bins(H, m1)
        fields = dict(kind=native.MODEL_FNO, constant_channels=self.constant_channels,
                      prescribed_channels=self.prescribed_channels, prognostic_channels=self.prognostic_channels,
                      context_size=self.context_size, fft_norm=native.FFT_NORM_FORWARD, n_cols=m2,
                      hidden_channels=self.hidden_channels, lifting_channels=self.lifting_channels,
                      projection_channels=self.projection_channels, n_layers=self.n_layers)
        return fields, rows_in, rows_out

    def _native_weights(self, H, W):
        m1, m2 = self._modes_for(H, W)
        rows = centre_crop(self._stored[0], m1)
        for name, p in self.fno.named_parameters():
            if ".convs.weight." in name:
                if name.endswith(".tensor"):
                    p = p[:, :, rows, :m2]
                elif name.endswith("factor_2"):
                    p = p[rows]
                elif name.endswith("factor_3"):
                    p = p[:m2]
            yield name, p


class TFNO2DModule(FNO2DModule):
    """``models/fno/fno.py:109-146``: Tucker-factorised spectral weights (``rank``).  The native
    library contracts the factors to the dense weight once per weight upload."""

    _factorization = "Tucker"

    def __init__(self, n_modes: list = [12, 12], constant_channels: int = 4, prescribed_channels: int = 1,
                 prognostic_channels: int = 8, hidden_channels: int = 32, lifting_channels: int = 256,
                 projection_channels: int = 256, n_layers: int = 4, max_n_modes: int = None, rank: float = 1.0,
                 bias: bool = True, context_size: int = 10, **kwargs):
        nn.Module.__init__(self)
        self._init_native()
        self.rank = rank
        self._configure(n_modes, constant_channels, prescribed_channels, prognostic_channels, hidden_channels,
                        lifting_channels, projection_channels, n_layers, max_n_modes, context_size, rank=rank)


# ======================================================================================
# FourCastNet (AFNONet) parameter containers, fourcastnet.py:41-57,60-76,156-178,530-537
# ======================================================================================
class _AFNO2DParams(nn.Module):
    def __init__(self, hidden_size, num_blocks, hidden_size_factor=1, scale=0.02):
        super().__init__()
        assert hidden_size % num_blocks == 0, \
            f"hidden_size {hidden_size} should be divisble by num_blocks {num_blocks}"
        bs = hidden_size // num_blocks
        self.w1 = nn.Parameter(scale * torch.randn(2, num_blocks, bs, bs * hidden_size_factor))
        self.b1 = nn.Parameter(scale * torch.randn(2, num_blocks, bs * hidden_size_factor))
        self.w2 = nn.Parameter(scale * torch.randn(2, num_blocks, bs * hidden_size_factor, bs))
        self.b2 = nn.Parameter(scale * torch.randn(2, num_blocks, bs))


class _MlpParams(nn.Module):
    def __init__(self, dim, hidden):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)


class _FNO2DFilterParams(nn.Module):
    """``fourcastnet.py:130-143``: the one-layer neuralop FNO behind ``filter: FNO2D`` (parameters under ``fno.``)."""

    n_modes = (12, 12)

    def __init__(self, hidden_size):
        super().__init__()
        stored = (self.n_modes[0], self.n_modes[1] // 2 + 1)
        self.fno = _FNOParams(hidden_size, hidden_size, hidden_size, hidden_size, hidden_size, 1, stored, None, 1.0)


class _BlockParams(nn.Module):
    def __init__(self, dim, mlp_hidden, num_blocks, filter="AFNO2D"):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, eps=1e-6)
        self.filter = _AFNO2DParams(dim, num_blocks) if filter == "AFNO2D" else _FNO2DFilterParams(dim)
        self.norm2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _MlpParams(dim, mlp_hidden)


class _PatchEmbedParams(nn.Module):
    def __init__(self, patch_size, in_chans, embed_dim):
        super().__init__()
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size)


class AFNONet(_NativeRollout):
    """``models/fourcastnet/fourcastnet.py:214-361`` with the ``AFNO2D`` filter (``:60-127``) or the ``FNO2D`` filter
    (``:130-153``: a one-layer neuralop FNO as the token mixer)."""

    _trainable = True   # training.afno_backbone

    def __init__(self, img_height: int = 720, img_width: int = 1440, patch_size: tuple = (16, 16),
                 constant_channels: int = 4, prescribed_channels: int = 0, prognostic_channels: int = 1,
                 filter: str = "AFNO2D", embed_dim: int = 768, depth: int = 12, mlp_ratio: float = 4.0,
                 drop_rate: float = 0.0, drop_path_rate: float = 0.0, num_blocks: int = 16,
                 sparsity_threshold: float = 0.01, hard_thresholding_fraction: float = 1.0, context_size: int = 1,
                 use_pos_embed: bool = True, **kwargs):
        super().__init__()
        self._init_native()
        if filter not in ("AFNO2D", "FNO2D"):
            raise NotImplementedError(f"filter={filter!r}: the AFNO2D and FNO2D filters of fourcastnet.py are implemented "
                                      "natively (SFNO needs torch_harmonics)")
        self.filter_kind = filter
        if drop_rate or drop_path_rate:
            raise NotImplementedError("forward-only implementation: drop_rate / drop_path_rate must be 0")
        self.img_size = (int(img_height), int(img_width))
        self.patch_size = (int(patch_size[0]), int(patch_size[1]))
        self.context_size = int(context_size)
        self.constant_channels = int(constant_channels)
        self.prescribed_channels = int(prescribed_channels)
        self.prognostic_channels = int(prognostic_channels)
        self.in_chans = constant_channels + (prescribed_channels + prognostic_channels) * context_size
        self.out_chans = prognostic_channels
        self.embed_dim = self.num_features = int(embed_dim)
        self.num_blocks, self.depth = int(num_blocks), int(depth)
        self.mlp_hidden = int(embed_dim * mlp_ratio)
        self.sparsity_threshold = float(sparsity_threshold)
        self.hard_thresholding_fraction = float(hard_thresholding_fraction)
        self.use_pos_embed = bool(use_pos_embed)
        if self.img_size[0] % self.patch_size[0] or self.img_size[1] % self.patch_size[1]:
            raise ValueError(f"patch_size {self.patch_size} does not divide img_size {self.img_size}")
        self.h = self.img_size[0] // self.patch_size[0]
        self.w = self.img_size[1] // self.patch_size[1]

        self.patch_embed = _PatchEmbedParams(self.patch_size, self.in_chans, self.embed_dim)
        if self.use_pos_embed:
            self.pos_embed = nn.Parameter(torch.zeros(1, self.h * self.w, self.embed_dim))
            nn.init.trunc_normal_(self.pos_embed, std=0.02)
        self.blocks = nn.ModuleList(_BlockParams(self.embed_dim, self.mlp_hidden, self.num_blocks, filter)
                                    for _ in range(self.depth))
        self.norm = nn.LayerNorm(self.embed_dim, eps=1e-6)   # held for state_dict parity; never applied (:263)
        self.head = nn.Linear(self.embed_dim, self.out_chans * self.patch_size[0] * self.patch_size[1], bias=False)
        for m in self.modules():                             # _init_weights, :270-277
            if isinstance(m, nn.Linear):
                nn.init.trunc_normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.LayerNorm):
                nn.init.constant_(m.bias, 0)
                nn.init.constant_(m.weight, 1.0)

    def _desc(self, H, W):
        if (H, W) != self.img_size:
            raise AssertionError(f"Input image size ({H}*{W}) doesn't match model ({self.img_size[0]}*{self.img_size[1]}).")
        if self.filter_kind == "FNO2D":
            m1, cols = self._fno2d_modes()
            rows, rows_out = neuralop_row_bins(self.h, m1)
            norm, filt = native.FFT_NORM_FORWARD, native.AFNO_FILTER_FNO2D
        else:
            rows, cols = afno_kept_region(self.h, self.w // 2 + 1, self.hard_thresholding_fraction)
            if not rows or cols <= 0:
                raise NotImplementedError("hard_thresholding_fraction keeps no Fourier mode")
            rows_out, norm, filt = rows, native.FFT_NORM_ORTHO, native.AFNO_FILTER_AFNO2D
        fields = dict(kind=native.MODEL_AFNO, afno_filter=filt, constant_channels=self.constant_channels,
                      prescribed_channels=self.prescribed_channels, prognostic_channels=self.prognostic_channels,
                      context_size=self.context_size, fft_norm=norm, n_cols=cols,
                      embed_dim=self.embed_dim, depth=self.depth, num_blocks=self.num_blocks,
                      mlp_hidden=self.mlp_hidden, patch_h=self.patch_size[0], patch_w=self.patch_size[1],
                      use_pos_embed=int(self.use_pos_embed), hidden_size_factor=1,
                      sparsity_threshold=self.sparsity_threshold, layernorm_eps=1e-6)
        return fields, rows, rows_out

    def _fno2d_modes(self):
        n1, n2 = _FNO2DFilterParams.n_modes
        return min(n1, self.h), min(n2 // 2 + 1, self.w // 2 + 1)

    def _native_weights(self, H, W):
        for name, p in self.named_parameters():
            if name.startswith("norm."):
                continue
            if name.endswith(".convs.weight.0.tensor"):      # FNO2D filter: the modes the token grid can hold
                m1, m2 = self._fno2d_modes()
                p = p[:, :, centre_crop(_FNO2DFilterParams.n_modes[0], m1), :m2]
            yield name, p


FourCastNet = AFNONet   # models/__init__.py:6 exports the class under this name
