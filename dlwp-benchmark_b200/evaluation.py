"""On-device evaluation tail: what ``scripts/evaluate.py`` does after the rollout, without the host-side numpy / xarray passes.

* ``denormalize_split``  -- ``evaluate.py:281-296``: undo the per-variable normalisation; the result is one contiguous
  ``[B, T, H, W]`` array per variable (one kernel: one read and one write of the forecast).
* ``write_forecast``     -- ``build_dataset`` (``evaluate.py:320-416``): one netCDF file with coordinates ``sample, time, lat,
  lon`` and one variable per prognostic field with its statistics as attributes.  Written as netCDF-3 (64-bit offset)
  through ``scipy.io`` (xarray / netCDF4 are not in this image; the file opens with ``xarray.open_dataset`` elsewhere).
  ``deg`` defaults to the grid's own resolution ``180 / H`` with cell-centre latitudes (upstream convention); the fork
  hard-codes ``deg = 1.0`` and ``lat = arange(-90, 90, deg)`` (``evaluate.py:344,353``): pass ``fork_coordinates=True`` for that.
* RMSE / ACC stay in ``metrics.py`` (``evaluate.py:786-827``).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import native


def variable_names(prognostic_variable_names_and_levels: Dict[str, Sequence]) -> list:
    """Variable order of the channel axis (``evaluate.py:283-296``, ``:366-383``): ``{name}{level}`` per level, else ``name``."""
    names = []
    for p, levels in prognostic_variable_names_and_levels.items():
        names += [f"{p}{l}" for l in levels] if len(levels) > 0 else [p]
    return names


def variable_stats(prognostic_variable_names_and_levels: Dict[str, Sequence], stats: Dict) -> list:
    """(mean, std) per channel from the dataset statistics dictionary (``dataset.stats`` of the reference)."""
    out = []
    for p, levels in prognostic_variable_names_and_levels.items():
        if len(levels) > 0:
            out += [(float(stats[p]["level"][l]["mean"]), float(stats[p]["level"][l]["std"])) for l in levels]
        else:
            out.append((float(stats[p]["mean"]), float(stats[p]["std"])))
    return out


def denormalize_split(x: torch.Tensor, mean=None, std=None) -> torch.Tensor:
    """``x [B, T, C, H, W]`` (CUDA) -> ``[C, B, T, H, W]`` with ``x * std[c] + mean[c]``."""
    if x.device.type != "cuda":
        raise native.NativeError("denormalize_split runs on CUDA tensors only (no CPU fallback)")
    B, T, Cn, H, W = x.shape
    x = x.to(torch.float32).contiguous()
    dev = x.device

    def vec(v):
        return None if v is None else torch.as_tensor(v, dtype=torch.float32).to(dev).contiguous()

    m, s = vec(mean), vec(std)
    out = torch.empty(Cn, B, T, H, W, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev):
        st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        native.check(native.lib().dlwp_denormalize_split(
            C.c_void_p(x.data_ptr()), C.c_void_p(m.data_ptr()) if m is not None else None,
            C.c_void_p(s.data_ptr()) if s is not None else None, C.c_void_p(out.data_ptr()), B, T, Cn, H, W, st),
            "dlwp_denormalize_split")
    return out


def grid_coordinates(H: int, W: int, fork_coordinates: bool = False):
    if fork_coordinates:      # evaluate.py:344,353-354 as the fork has them
        deg = 1.0
        return np.arange(-90, 90, deg, dtype=np.float32)[:H], np.arange(0, 360, deg, dtype=np.float32)[:W]
    dlat, dlon = 180.0 / H, 360.0 / W
    return (np.arange(H, dtype=np.float32) * dlat - 90.0 + dlat / 2).astype(np.float32), (np.arange(W, dtype=np.float32) * dlon).astype(np.float32)


def write_forecast(path: str, per_variable: torch.Tensor, names: Sequence[str], init_hours: Optional[Sequence[float]] = None,
                   step_hours: float = 6.0, attrs: Optional[Sequence[Dict]] = None, fork_coordinates: bool = False):
    """``per_variable [C, B, T, H, W]`` (the output of ``denormalize_split``; a ``[C, B, H, W]`` tensor writes the
    initial conditions file) -> one netCDF-3 file, dims ``sample, (time,) lat, lon``, one variable per name."""
    from scipy.io import netcdf_file
    a = per_variable.detach().cpu().numpy() if isinstance(per_variable, torch.Tensor) else np.asarray(per_variable)
    with_time = a.ndim == 5
    Cn, B = a.shape[0], a.shape[1]
    H, W = a.shape[-2], a.shape[-1]
    assert len(names) == Cn, "one name per variable"
    lat, lon = grid_coordinates(H, W, fork_coordinates)
    with netcdf_file(path, "w", version=2) as f:
        f.createDimension("sample", B)
        if with_time:
            f.createDimension("time", a.shape[2])
        f.createDimension("lat", H)
        f.createDimension("lon", W)
        v = f.createVariable("sample", "f8", ("sample",))
        v[:] = np.arange(B, dtype=np.float64) if init_hours is None else np.asarray(init_hours, dtype=np.float64)
        v.units = "hours since 1970-01-01 00:00:00"
        if with_time:
            v = f.createVariable("time", "f8", ("time",))
            v[:] = step_hours * np.arange(1, a.shape[2] + 1, dtype=np.float64)      # lead times dt, 2 dt, ... (evaluate.py:345-346)
            v.units = "hours"
        v = f.createVariable("lat", "f4", ("lat",))
        v[:] = lat
        v.units = "degrees_north"
        v = f.createVariable("lon", "f4", ("lon",))
        v[:] = lon
        v.units = "degrees_east"
        dims = ("sample", "time", "lat", "lon") if with_time else ("sample", "lat", "lon")
        for c, name in enumerate(names):
            v = f.createVariable(name, "f4", dims)
            v[:] = a[c]
            if attrs is not None:
                for k, val in attrs[c].items():
                    if isinstance(val, (int, float, str)):
                        setattr(v, k, val)
    return path
