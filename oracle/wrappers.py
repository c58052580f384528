"""Oracle: rollout loop and the ``FNO2DModule`` / ``TFNO2DModule`` wrappers.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows ``/root/reference/src/dlwpbench/models/fno/fno.py``: channel count :36,
``_prepare_inputs`` :49-62, the autoregressive loop :78-106, TFNO construction :131-146.
Two deliberate deviations, both documented in DESIGN.md:

* the fork's context-blending line calls ``.to()`` on a Python list (fno.py:91-95,
  fourcastnet.py:336-340) and crashes for more than one step; the loop below is the
  working upstream form that the same repo keeps at fourcastnet.py:502-506;
* the fork moves every step's output to the host (fno.py:104); outputs stay on the
  input device here (as fourcastnet.py:525-527 does).
"""

from __future__ import annotations

import torch
import torch.nn as nn

from .neuralop_fno import FNO, TFNO


def prepare_inputs(constants=None, prescribed=None, prognostic=None):
    """constants ``[B,1,Cc,H,W]`` | prescribed ``[B,ctx,Cp,H,W]`` | prognostic ``[B,ctx,Cg,H,W]``
    -> ``[B, Cc + ctx*Cp + ctx*Cg, H, W]`` (time-major inside each group)."""
    parts = []
    if constants is not None:
        parts.append(constants[:, 0])
    if prescribed is not None:
        parts.append(prescribed.flatten(1, 2))
    if prognostic is not None:
        parts.append(prognostic.flatten(1, 2))
    return torch.cat(parts, dim=1)


def rollout(backbone, context_size, constants=None, prescribed=None, prognostic=None):
    """``out_t = prognostic_t[:, -1] + backbone(x_t)`` for ``t = ctx .. T-1``; returns
    ``[B, T-ctx, Cg, H, W]`` on the input device."""
    ctx = context_size
    outs = []
    for t in range(ctx, prognostic.shape[1]):
        if t == ctx:
            prog_t = prognostic[:, 0:ctx]
        else:
            own = torch.stack(outs, dim=1)[:, -ctx:]
            prog_t = torch.cat([prognostic[:, t - ctx:ctx], own], dim=1) if t - ctx < ctx else own
        presc_t = prescribed[:, t - ctx:t] if prescribed is not None else None
        x_t = prepare_inputs(constants, presc_t, prog_t)
        outs.append(prog_t[:, -1] + backbone(x_t))
    return torch.stack(outs, dim=1)


class FNO2DModule(nn.Module):
    def __init__(self, n_modes=(12, 12), constant_channels=4, prescribed_channels=1, prognostic_channels=8,
                 hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4,
                 max_n_modes=None, bias=True, context_size=10, **kwargs):
        super().__init__()
        self.context_size = context_size
        in_channels = constant_channels + (prescribed_channels + prognostic_channels) * context_size
        self.fno = self._build(n_modes=list(n_modes), max_n_modes=max_n_modes, in_channels=in_channels,
                               hidden_channels=hidden_channels, lifting_channels=lifting_channels,
                               projection_channels=projection_channels, out_channels=prognostic_channels,
                               n_layers=n_layers, **self._extra(kwargs))

    @staticmethod
    def _build(**kw):
        return FNO(**kw)

    @staticmethod
    def _extra(kwargs):
        return {}

    def forward(self, constants=None, prescribed=None, prognostic=None):
        return rollout(self.fno, self.context_size, constants, prescribed, prognostic)


class TFNO2DModule(FNO2DModule):
    def __init__(self, *args, rank=1.0, **kwargs):
        self._rank = rank
        super().__init__(*args, rank=rank, **kwargs)

    @staticmethod
    def _build(**kw):
        return TFNO(**kw)

    @staticmethod
    def _extra(kwargs):
        return {"rank": kwargs.get("rank", 1.0)}
