"""B200-native rollout forward pass of dlwp-benchmark's FNO2D / TFNO2D / FourCastNet-AFNO.

``from dlwp_benchmark_b200 import FNO2DModule, TFNO2DModule, FourCastNet`` gives drop-in
replacements of the reference's model classes (see ``models.py``); the compute lives in
``libdlwp_b200.so`` (C ABI: ``include/dlwp_b200.h``) and there is no CPU fallback.
"""
from . import native                                             # noqa: F401
from .models import AFNONet, FNO2DModule, FourCastNet, TFNO2DModule  # noqa: F401
from .metrics import lat_weights, rollout_metrics                # noqa: F401
from .spectral import FourierResidualBlock, SpectralConv2d       # noqa: F401
from .prefetch import DevicePrefetcher                           # noqa: F401

__all__ = ["FNO2DModule", "TFNO2DModule", "AFNONet", "FourCastNet", "lat_weights", "rollout_metrics", "native",
           "SpectralConv2d", "FourierResidualBlock", "DevicePrefetcher"]
