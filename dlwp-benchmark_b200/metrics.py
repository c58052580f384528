"""Latitude-weighted RMSE / ACC per lead time and variable, accumulated on the device.

Replaces the host-side xarray computation of ``scripts/evaluate.py:786-827``: weights
``cos(lat) / mean(cos(lat))`` (:790-792), RMSE (:801-802), ACC (:814-820).  Each rank
accumulates fp64 partial sums ``[T, C, 4]`` over its shard of the batch / ensemble with one
kernel launch; the only inter-GPU traffic of the whole rollout is one all-reduce of that
array (NCCL when the tensors are on CUDA).
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import native


def lat_weights(H: int, device=None) -> torch.Tensor:
    """``cos(lat) / mean(cos(lat))`` on the equiangular cell-centre grid with ``H`` rows."""
    d = 180.0 / H
    lat = torch.arange(H, dtype=torch.float64) * d + (-90.0 + d / 2)
    c = torch.cos(lat * (math.pi / 180.0))
    return (c / c.mean()).to(torch.float32).to(device)


def partial_sums(outputs, targets, climatology=None, weights=None, out=None) -> torch.Tensor:
    """``[T, C, 4]`` fp64 sums over (sample, lat, lon) of ``w (o-t)^2, w (o-c)(t-c), w (o-c)^2,
    w (t-c)^2`` for device tensors ``[B, T, C, H, W]``.  ``out``: a zeroed ``[T, C, 4]`` fp64 device
    tensor to accumulate into (a view of a larger reduction buffer, see ``rollout_metrics``)."""
    if outputs.device.type != "cuda":
        raise native.NativeError("metrics run on CUDA tensors only (no CPU fallback)")
    B, T, Cn, H, W = outputs.shape
    o = outputs.to(torch.float32).contiguous()
    t = targets.to(device=o.device, dtype=torch.float32).contiguous()
    w = (lat_weights(H, o.device) if weights is None else weights.to(device=o.device, dtype=torch.float32)).contiguous()
    c = None if climatology is None else climatology.to(device=o.device, dtype=torch.float32).contiguous()
    sums = torch.zeros(T, Cn, 4, dtype=torch.float64, device=o.device) if out is None else out
    assert sums.shape == (T, Cn, 4) and sums.dtype == torch.float64 and sums.is_contiguous()
    with torch.cuda.device(o.device):
        st = C.c_void_p(torch.cuda.current_stream(o.device).cuda_stream)
        native.check(native.lib().dlwp_metrics_accumulate(
            C.c_void_p(o.data_ptr()), C.c_void_p(t.data_ptr()), C.c_void_p(c.data_ptr()) if c is not None else None,
            C.c_void_p(w.data_ptr()), B, T, Cn, H, W, C.c_void_p(sums.data_ptr()), st), "dlwp_metrics_accumulate")
    return sums


def finalize(sums: torch.Tensor, count):
    """Reduced sums ``[T, C, 4]`` and the number of (sample, lat, lon) points (a number, or a 0-d tensor
    on the sums' device) -> (rmse, acc)."""
    mean = sums / count
    rmse = torch.sqrt(mean[..., 0])
    acc = mean[..., 1] / torch.sqrt(mean[..., 2] * mean[..., 3])
    return rmse, acc


def reduce_sums(sums: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the per-rank partials over the process group IN PLACE: the rollout's only collective (one
    all-reduce of a few KB; nothing is copied from or to the host and nothing synchronises)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    return sums


def rollout_metrics(outputs, targets, climatology=None, weights=None, group=None, total_samples=None):
    """(rmse ``[T, C]``, acc ``[T, C]``) over the whole (possibly sharded) ensemble.  ``total_samples``: the
    number of members over all ranks when the caller knows it (an even split does); otherwise the per-rank
    counts ride in the same all-reduce as one extra element."""
    B, T, Cn, H, W = outputs.shape
    buf = torch.zeros(T * Cn * 4 + 1, dtype=torch.float64, device=outputs.device)
    sums = buf[:-1].view(T, Cn, 4)
    partial_sums(outputs, targets, climatology, weights, out=sums)
    if total_samples is not None:
        reduce_sums(sums, group)
        return finalize(sums, float(total_samples) * H * W)
    buf[-1:].fill_(float(B * H * W))
    reduce_sums(buf, group)
    return finalize(sums, buf[-1])
