"""Oracle helper: import the REAL reference source files, unmodified, under import stubs.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Works only where
``/root/reference`` exists (the build container); the GPU box never has it, so nothing
run there may call this.  Used by ``tests/golden/make_golden.py`` and by the CPU tests that
pin ``oracle/afno.py`` and ``oracle/spectral2d.py`` against the reference itself.

Stubs (all for names the AFNO / SpectralConv2d arithmetic never executes):
``neuralop.models.{FNO,TFNO}``, ``torch_harmonics.examples.sfno``, ``timm.models.layers``
(``DropPath`` -> identity, ``trunc_normal_`` -> ``torch.nn.init.trunc_normal_``) and
``numpy.lib.arraypad.pad`` (removed in numpy 2).
"""

from __future__ import annotations

import importlib.util
import os
import sys
import types

REF_ROOT = "/root/reference/src/dlwpbench"


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "models/fourcastnet/fourcastnet.py"))


def _stub(name, **attrs):
    if name in sys.modules:
        mod = sys.modules[name]
    else:
        mod = types.ModuleType(name)
        mod.__dict__["__graft_stub__"] = True
        sys.modules[name] = mod
    for k, v in attrs.items():
        if not hasattr(mod, k):
            setattr(mod, k, v)
    return mod


def _install_stubs():
    import numpy as np
    import torch.nn as nn

    class _Absent(nn.Module):
        def __init__(self, *a, **k):
            raise RuntimeError("third-party model is not available in this image")

    _stub("neuralop")
    _stub("neuralop.models", FNO=_Absent, TFNO=_Absent)
    _stub("torch_harmonics")
    _stub("torch_harmonics.examples")
    _stub("torch_harmonics.examples.sfno", SphericalFourierNeuralOperatorNet=_Absent)
    _stub("timm")
    _stub("timm.models")
    _stub("timm.models.layers", DropPath=nn.Identity, trunc_normal_=nn.init.trunc_normal_)
    try:
        import numpy.lib.arraypad as ap  # noqa: F401
        if not hasattr(ap, "pad"):
            ap.pad = np.pad
    except Exception:
        _stub("numpy.lib.arraypad", pad=np.pad)


def load_fourcastnet():
    """The reference ``models/fourcastnet/fourcastnet.py`` as a module object."""
    _install_stubs()
    path = os.path.join(REF_ROOT, "models/fourcastnet/fourcastnet.py")
    spec = importlib.util.spec_from_file_location("_ref_fourcastnet", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_unet_spectral():
    """``batchmul2d`` / ``SpectralConv2d`` / ``FourierResidualBlock`` executed from the reference
    file's own text (unet.py:14-69 and :780-836; the file's top-level imports need hydra)."""
    import torch as th
    path = os.path.join(REF_ROOT, "models/unet/unet.py")
    lines = open(path).read().split("\n")
    ns = {"th": th, "__name__": "_ref_unet_slice"}
    exec(compile("\n".join(lines[13:69]), path + ":14-69", "exec"), ns)
    exec(compile("\n".join(lines[779:836]), path + ":780-836", "exec"), ns)
    return types.SimpleNamespace(**{k: ns[k] for k in ("batchmul2d", "SpectralConv2d", "FourierResidualBlock")})


def reference_afno_backbone(model, x):
    """One evaluation of the reference ``AFNONet`` on a prepared input ``[B,Cin,H,W]``, i.e.
    fourcastnet.py:345-357 driven from outside (its own multi-step loop crashes, :336-340)."""
    from einops import rearrange
    t = model.head(model.forward_features(x))
    return rearrange(t, "b h w (p1 p2 c_out) -> b c_out (h p1) (w p2)", p1=model.patch_size[0],
                     p2=model.patch_size[1], h=model.img_size[0] // model.patch_size[0],
                     w=model.img_size[1] // model.patch_size[1])
