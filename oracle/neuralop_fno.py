"""Oracle: from-spec restatement of ``neuralop`` 0.3.x ``FNO`` / ``TFNO`` (2-D case).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  **PARITY UNPINNED**: the
arithmetic lives in the third-party package ``neuraloperator`` (import name
``neuralop``, git 05c01c3 per ``/root/reference/README.md:29-39``) with ``tltorch`` /
``tensorly`` for the Tucker weights; none of them is vendored under ``/root/reference``
or installed in this image and there is no network.  What anchors this file:

* the call sites ``models/fno/fno.py:38-47`` (FNO) and ``:136-146`` (TFNO, ``rank=``):
  only ``n_modes, max_n_modes, in/hidden/lifting/projection/out channels, n_layers``
  (+ ``rank``) are passed, so every other knob is the neuralop default:
  ``non_linearity=gelu`` (exact erf), ``use_mlp=False``, ``norm=None``,
  ``preactivation=False``, ``fno_skip="linear"``, ``separable=False``,
  ``fft_norm="forward"``, ``implementation="factorized"``, no domain padding;
* the in-repo dense contraction ``einsum("bixy,ioxy->boxy")`` of
  ``models/unet/unet.py:15-17``;
* the naming cross-check ``fno2d64m_cyl_d307``: ``307**2 * 12*7 * 2 * 4 = 63.3 M``
  parameters (``scripts/plot_results.py:27``), which pins ``n_modes[-1] // 2 + 1``.

Parameter names follow neuralop so a reference-shaped ``state_dict`` loads:
``lifting.fcs.{0,1}.{weight,bias}``, ``fno_blocks.fno_skips.{l}.weight``,
``fno_blocks.convs.weight.{l}.tensor`` (dense, real view ``[..., 2]``) or
``...weight.{l}.core`` + ``...weight.{l}.factors.factor_{k}`` (Tucker, real views),
``fno_blocks.convs.bias``, ``projection.fcs.{0,1}.{weight,bias}``.
"""

from __future__ import annotations

import math
from typing import List, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# neuralop.layers.mlp.MLP  (lifting / projection), SURVEY.md §8a row N5
# --------------------------------------------------------------------------------------
class PointwiseMLP(nn.Module):
    """``n_layers`` 1x1 convolutions with GELU between them (bias on every layer)."""

    def __init__(self, in_channels: int, out_channels: int, hidden_channels: int, n_layers: int = 2):
        super().__init__()
        self.n_layers = n_layers
        widths = [in_channels] + [hidden_channels] * (n_layers - 1) + [out_channels]
        self.fcs = nn.ModuleList(nn.Conv2d(widths[i], widths[i + 1], 1) for i in range(n_layers))

    def forward(self, x):
        for i, fc in enumerate(self.fcs):
            x = fc(x)
            if i < self.n_layers - 1:
                x = F.gelu(x)
        return x


# --------------------------------------------------------------------------------------
# Spectral weights: dense complex tensor, or complex Tucker (tltorch), row N2 / N4
# --------------------------------------------------------------------------------------
class DenseSpectralWeight(nn.Module):
    """Complex weight ``[Ci, Co, m1, m2']`` stored as a real view (tltorch ``ComplexDenseTensor``)."""

    def __init__(self, shape: Sequence[int], std: float):
        super().__init__()
        self.shape = tuple(shape)
        w = torch.randn(*shape, dtype=torch.cfloat) * std
        self.tensor = nn.Parameter(torch.view_as_real(w).clone())

    def to_dense(self) -> torch.Tensor:
        return torch.view_as_complex(self.tensor.contiguous())

    def contract(self, x: torch.Tensor) -> torch.Tensor:
        return torch.einsum("bixy,ioxy->boxy", x, self.to_dense().to(x.dtype))


def tucker_rank_from_fraction(shape: Sequence[int], rank: float) -> List[int]:
    """tensorly ``validate_tucker_rank`` (rounding='round', no fixed modes): find the
    per-mode fraction ``f`` such that core + factors hold ``rank * prod(shape)`` values."""
    shape = [int(s) for s in shape]
    n_dense = float(math.prod(shape))
    sq = float(sum(s * s for s in shape))
    order = len(shape)

    def fun(f):
        return n_dense * f ** order + sq * f - rank * n_dense

    lo, hi = 0.0, max(rank, 1.0)
    for _ in range(200):  # bisection; brentq in tensorly, same root
        mid = 0.5 * (lo + hi)
        if fun(mid) > 0:
            hi = mid
        else:
            lo = mid
    f = 0.5 * (lo + hi)
    return [max(int(round(s * f)), 1) for s in shape]


class FactorList(nn.Module):
    """Mirror of tltorch's ``FactorList`` naming: parameters ``factor_0 .. factor_{n-1}``."""

    def __init__(self, factors):
        super().__init__()
        for i, f in enumerate(factors):
            self.register_parameter(f"factor_{i}", nn.Parameter(f))
        self._n = len(factors)

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        return getattr(self, f"factor_{i}")


class TuckerSpectralWeight(nn.Module):
    """Complex Tucker weight: ``W = core x1 U_i x2 U_o x3 U_x x4 U_y`` (tltorch
    ``ComplexTuckerTensor``), parameters stored as real views ``[..., 2]``."""

    def __init__(self, shape: Sequence[int], rank, std: float):
        super().__init__()
        self.shape = tuple(shape)
        if isinstance(rank, float):
            ranks = tucker_rank_from_fraction(shape, rank)
        elif isinstance(rank, int):
            ranks = [min(rank, s) for s in shape]
        else:
            ranks = list(rank)
        self.rank = tuple(ranks)
        # tltorch TuckerTensor.normal_: std split evenly over core and factors
        r = math.prod(math.sqrt(v) for v in ranks)
        std_f = (std / r) ** (1.0 / (len(shape) + 1))
        core = torch.randn(*ranks, dtype=torch.cfloat) * std_f
        self.core = nn.Parameter(torch.view_as_real(core).clone())
        self.factors = FactorList(
            [torch.view_as_real(torch.randn(s, r_, dtype=torch.cfloat) * std_f).clone() for s, r_ in zip(shape, ranks)]
        )

    def _cfactors(self, dtype=None):
        out = []
        for k in range(len(self.factors)):
            f = torch.view_as_complex(self.factors[k].contiguous())
            out.append(f if dtype is None else f.to(dtype))
        return out

    def to_dense(self, dtype=None) -> torch.Tensor:
        core = torch.view_as_complex(self.core.contiguous())
        if dtype is not None:
            core = core.to(dtype)
        ui, uo, ux, uy = self._cfactors(dtype)
        return torch.einsum("abcd,ia,ob,xc,yd->ioxy", core, ui, uo, ux, uy)

    def contract(self, x: torch.Tensor) -> torch.Tensor:
        # neuralop _contract_tucker ("factorized" implementation): one einsum
        # "bixy,acde,ia,oc,xd,ye->boxy" over x, core and the four factor matrices, no dense
        # reconstruction.  torch.einsum without opt_einsum contracts left to right, which makes
        # that single call ~1e4 times slower than a pairwise order; the same sum is evaluated
        # here pairwise (input projection, weight-side products, mode-wise contraction), the
        # order an optimised contraction path takes.  Only the rounding order differs.
        core = torch.view_as_complex(self.core.contiguous()).to(x.dtype)
        ui, uo, ux, uy = self._cfactors(x.dtype)
        xa = torch.einsum("bixy,ia->baxy", x, ui)
        wa = torch.einsum("acde,oc->aode", core, uo)
        wa = torch.einsum("aode,xd->aoxe", wa, ux)
        wa = torch.einsum("aoxe,ye->aoxy", wa, uy)
        return torch.einsum("baxy,aoxy->boxy", xa, wa)


# --------------------------------------------------------------------------------------
# neuralop.layers.spectral_convolution.SpectralConv (0.3.x, fftshift formulation)
# --------------------------------------------------------------------------------------
def _centre_slices(sizes: Sequence[int], keep: Sequence[int]):
    """Slices neuralop builds: centred crop on every dim but the last, leading crop on the last."""
    starts = [s - min(s, k) for s, k in zip(sizes, keep)]
    sl = [slice(None), slice(None)]
    for st in starts[:-1]:
        sl.append(slice(st // 2, -st // 2) if st else slice(st, None))
    sl.append(slice(None, -starts[-1]) if starts[-1] else slice(None))
    return tuple(sl)


class SpectralConvLayers(nn.Module):
    """The ``n_layers`` spectral convolutions of one FNO, sharing a ``[n_layers, C, 1, 1]`` bias."""

    def __init__(self, in_channels, out_channels, n_modes, n_layers, max_n_modes=None,
                 factorization=None, rank=1.0, fft_norm="forward"):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        n_modes = list(n_modes)
        n_modes[-1] = n_modes[-1] // 2 + 1           # rfft redundancy folded into the stored count
        self.n_modes = n_modes
        if max_n_modes is None:
            max_n_modes = list(n_modes)
        else:
            max_n_modes = list(max_n_modes)
            max_n_modes[-1] = max_n_modes[-1] // 2 + 1
        self.max_n_modes = max_n_modes
        self.fft_norm = fft_norm
        std = (2.0 / (in_channels + out_channels)) ** 0.5
        shape = (in_channels, out_channels, *max_n_modes)
        if factorization is None:
            self.weight = nn.ModuleList(DenseSpectralWeight(shape, std) for _ in range(n_layers))
        elif factorization.lower() in ("tucker", "complextucker"):
            self.weight = nn.ModuleList(TuckerSpectralWeight(shape, rank, std) for _ in range(n_layers))
        else:
            raise ValueError(f"factorization {factorization!r} is not on the reference's path")
        self.bias = nn.Parameter(std * torch.randn(n_layers, out_channels, 1, 1))

    def dense_weight(self, index: int, dtype=None) -> torch.Tensor:
        w = self.weight[index]
        full = w.to_dense(dtype) if isinstance(w, TuckerSpectralWeight) else w.to_dense()
        if dtype is not None:
            full = full.to(dtype)
        return full

    def forward(self, x: torch.Tensor, index: int = 0) -> torch.Tensor:
        B, _, H, W = x.shape
        fft_size = [H, W // 2 + 1]
        xf = torch.fft.rfftn(x, norm=self.fft_norm, dim=(-2, -1))
        xf = torch.fft.fftshift(xf, dim=(-2,))
        out = torch.zeros(B, self.out_channels, *fft_size, dtype=xf.dtype, device=x.device)

        # weight is first cropped from max_n_modes to n_modes, then x to the weight's size
        w_sl = _centre_slices(self.max_n_modes, [min(s, n) for s, n in zip(fft_size, self.n_modes)])
        wmod = self.weight[index]
        if w_sl != _centre_slices(self.max_n_modes, self.max_n_modes):
            dense = self.dense_weight(index, xf.dtype)[w_sl]
            x_sl = _centre_slices(fft_size, dense.shape[2:])
            out[x_sl] = torch.einsum("bixy,ioxy->boxy", xf[x_sl], dense)
        else:
            x_sl = _centre_slices(fft_size, self.max_n_modes)
            out[x_sl] = wmod.contract(xf[x_sl])

        # NB: neuralop applies fftshift (not ifftshift) a second time; for odd H that is a
        # one-bin displacement, reproduced here because this is a restatement, not a fix.
        out = torch.fft.fftshift(out, dim=(-2,))
        y = torch.fft.irfftn(out, s=(H, W), dim=(-2, -1), norm=self.fft_norm)
        return y + self.bias[index]


class FNOBlocks(nn.Module):
    """Post-activation FNO blocks without MLP / norm (neuralop defaults), row N3."""

    def __init__(self, hidden_channels, n_modes, n_layers, max_n_modes=None, factorization=None, rank=1.0):
        super().__init__()
        self.n_layers = n_layers
        self.convs = SpectralConvLayers(hidden_channels, hidden_channels, n_modes, n_layers,
                                        max_n_modes=max_n_modes, factorization=factorization, rank=rank)
        self.fno_skips = nn.ModuleList(
            nn.Conv2d(hidden_channels, hidden_channels, 1, bias=False) for _ in range(n_layers))

    def forward(self, x, index: int):
        x = self.convs(x, index) + self.fno_skips[index](x)
        if index < self.n_layers - 1:
            x = F.gelu(x)
        return x


class FNO(nn.Module):
    """``neuralop.models.FNO`` with the keyword surface the reference uses (row N1)."""

    def __init__(self, n_modes, hidden_channels, in_channels=3, out_channels=1, lifting_channels=256,
                 projection_channels=256, n_layers=4, max_n_modes=None, factorization=None, rank=1.0, **_):
        super().__init__()
        self.n_modes = list(n_modes)
        self.n_layers = n_layers
        self.hidden_channels = hidden_channels
        self.in_channels, self.out_channels = in_channels, out_channels
        self.fno_blocks = FNOBlocks(hidden_channels, n_modes, n_layers, max_n_modes=max_n_modes,
                                    factorization=factorization, rank=rank)
        if lifting_channels:
            self.lifting = PointwiseMLP(in_channels, hidden_channels, lifting_channels, n_layers=2)
        else:
            self.lifting = PointwiseMLP(in_channels, hidden_channels, hidden_channels, n_layers=1)
        self.projection = PointwiseMLP(hidden_channels, out_channels, projection_channels, n_layers=2)

    def forward(self, x):
        x = self.lifting(x)
        for l in range(self.n_layers):
            x = self.fno_blocks(x, l)
        return self.projection(x)


class TFNO(FNO):
    """``neuralop.models.TFNO`` = ``FNO(factorization="Tucker")``."""

    def __init__(self, *args, rank=1.0, **kwargs):
        kwargs.pop("factorization", None)
        super().__init__(*args, factorization="Tucker", rank=rank, **kwargs)


# --------------------------------------------------------------------------------------
# Index bookkeeping shared with the tests: which FFT row bins the layer reads and writes.
# --------------------------------------------------------------------------------------
def neuralop_row_bins(H: int, m1: int):
    """Return ``(bins_in, bins_out)``: for weight row ``j`` the unshifted FFT row bin that
    is read, and the bin its product is written to, under the double-``fftshift`` form."""
    m = min(H, m1)
    sl = _centre_slices([H, 1], [m, 1])[2]
    shifted = torch.arange(H)[sl]                        # positions in the shifted spectrum
    # fftshift: out[i] = in[(i - H//2) mod H]  (roll by +H//2)
    bins_in = (shifted - H // 2) % H                      # bin that sits at shifted position i
    # second fftshift on out_fft: value at position i moves to (i + H//2) mod H
    bins_out = (shifted + H // 2) % H
    return bins_in.tolist(), bins_out.tolist()
