"""Oracle: latitude-weighted RMSE and ACC per lead time and variable.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows ``/root/reference/src/dlwpbench/scripts/evaluate.py``: weights :790-792
(``cos(lat) / mean(cos(lat))``), RMSE :801-802, ACC :814-820 (mean over sample, lat, lon).
The reference evaluates them in xarray on the host; here they are plain torch in fp64.
"""

from __future__ import annotations

import math

import torch


def latitudes(H: int) -> torch.Tensor:
    """Cell-centre latitudes of an equiangular grid with ``H`` rows, in degrees (WeatherBench
    convention ``arange(-90 + d/2, 90, d)`` with ``d = 180 / H``)."""
    d = 180.0 / H
    return torch.arange(H, dtype=torch.float64) * d + (-90.0 + d / 2)


def lat_weights(H: int) -> torch.Tensor:
    c = torch.cos(latitudes(H) * (math.pi / 180.0))
    return c / c.mean()


def partial_sums(outputs, targets, climatology=None, weights=None):
    """Per-(lead time, variable) sums ``[T, C, 4]`` over (sample, lat, lon) in fp64:
    ``sum w*(o-t)^2, sum w*(o-c)*(t-c), sum w*(o-c)^2, sum w*(t-c)^2`` and the element count.
    ``outputs``/``targets``: ``[B, T, C, H, W]``; ``climatology``: ``[C, H, W]`` or ``None`` (= 0)."""
    o, t = outputs.double(), targets.double()
    B, T, C, H, W = o.shape
    w = (lat_weights(H) if weights is None else weights.double()).view(1, 1, 1, H, 1)
    c = torch.zeros(C, H, W, dtype=torch.float64) if climatology is None else climatology.double()
    c = c.view(1, 1, C, H, W)
    s = torch.stack([
        (w * (o - t) ** 2).sum(dim=(0, 3, 4)),
        (w * (o - c) * (t - c)).sum(dim=(0, 3, 4)),
        (w * (o - c) ** 2).sum(dim=(0, 3, 4)),
        (w * (t - c) ** 2).sum(dim=(0, 3, 4)),
    ], dim=-1)
    return s, B * H * W


def finalize(sums: torch.Tensor, count: int):
    """``sums [T, C, 4]`` (already reduced over ranks) -> ``(rmse [T, C], acc [T, C])``."""
    mean = sums / float(count)
    rmse = torch.sqrt(mean[..., 0])
    acc = mean[..., 1] / torch.sqrt(mean[..., 2] * mean[..., 3])
    return rmse, acc


def rmse_acc(outputs, targets, climatology=None, weights=None):
    s, n = partial_sums(outputs, targets, climatology, weights)
    return finalize(s, n)
