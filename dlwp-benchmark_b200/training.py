"""Training-mode forward / backward of ``FNO2DModule`` / ``TFNO2DModule``: ``scripts/train.py:263-271`` calls
``model(...)``, a loss and ``.backward()``.  With grad mode on, the modules route here instead of the inference plan:

* every FLOP of the backbone runs in ``libdlwp_b200.so`` (fp32 kernel set): ``dlwp_linear_forward/_backward`` (neuralop MLP
  1x1 convolutions, FNO skip convolution + bias + GELU epilogue) and ``dlwp_spectral_forward/_backward`` (dense Fourier layer
  and its adjoint + per-mode weight gradient), wrapped in ``torch.autograd.Function`` objects on channels-last tensors;
* torch autograd only glues: the step-input concatenation and residual frame of the rollout (``fno.py:49-62,78-106``) and,
  for TFNO, the reconstruction of the dense spectral weight from the Tucker core and factors (a 86 k-entry einsum whose
  gradient autograd carries back to ``core`` / ``factors``).

FourCastNet (AFNO) has no training path yet and keeps raising ``NotImplementedError`` under grad mode.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import native
from .plan import centre_crop, neuralop_row_bins


def _st(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


class _Layout(torch.autograd.Function):
    """NCHW ``[B, C, H, W]`` <-> channels-last ``[B*H*W, C]`` (native transposes; the backward is the other direction)."""

    @staticmethod
    def forward(ctx, x, to_last, shape):
        B, Cn, H, W = shape
        ctx.to_last, ctx.shape = to_last, shape
        x = x.contiguous()
        out = torch.empty((B * H * W, Cn) if to_last else (B, Cn, H, W), device=x.device, dtype=torch.float32)
        fn = native.lib().dlwp_nchw_to_nhwc if to_last else native.lib().dlwp_nhwc_to_nchw
        native.check(fn(_p(x), _p(out), B, Cn, H * W, _st(x.device)), "layout")
        return out

    @staticmethod
    def backward(ctx, g):
        return _Layout.apply(g, not ctx.to_last, ctx.shape), None, None


class _Linear(torch.autograd.Function):
    """``y = act(x W^T + b + pre)`` on ``[P, K]`` rows; ``W`` in the Conv2d / Linear layout ``[N, K(, 1, 1)]``."""

    @staticmethod
    def forward(ctx, x, W, b, pre, act):
        x = x.contiguous()
        P, K = x.shape
        N = W.shape[0]
        W2 = W.reshape(N, K).contiguous()
        y = torch.empty(P, N, device=x.device, dtype=torch.float32)
        z = torch.empty_like(y) if act else None
        native.check(native.lib().dlwp_linear_forward(_p(x), _p(W2), _p(b), _p(pre.contiguous() if pre is not None else None), _p(y),
                                                      _p(z), P, K, N, int(act), _st(x.device)), "dlwp_linear_forward")
        ctx.save_for_backward(x, W2, z)
        ctx.act, ctx.wshape, ctx.has_b, ctx.has_pre = int(act), W.shape, b is not None, pre is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, W2, z = ctx.saved_tensors
        dy = dy.contiguous()
        P, K = x.shape
        N = W2.shape[0]
        dz = torch.empty(P, N, device=x.device, dtype=torch.float32)
        dx = torch.empty(P, K, device=x.device, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        dW = torch.empty(N, K, device=x.device, dtype=torch.float32)
        db = torch.empty(N, device=x.device, dtype=torch.float32) if ctx.has_b else None
        wt = torch.empty(K, N, device=x.device, dtype=torch.float32) if dx is not None else None
        native.check(native.lib().dlwp_linear_backward(_p(dy), _p(z), _p(x), _p(W2), _p(dz), _p(dx), _p(dW), _p(db), _p(wt), P, K, N,
                                                       ctx.act, _st(x.device)), "dlwp_linear_backward")
        return dx, dW.reshape(ctx.wshape), db, (dz if ctx.has_pre else None), None


class _Spectral(torch.autograd.Function):
    """Dense Fourier layer on channels-last ``x [B*H*W, Ci]`` with the complex weight as a real view ``[Ci, Co, R, Kc, 2]``."""

    @staticmethod
    def forward(ctx, x, weight, geom):
        B, H, W, rows_in, rows_out, fft_norm = geom
        x = x.contiguous()
        weight = weight.contiguous()
        Ci, Co, R, Kc = weight.shape[:4]
        lib = native.lib()
        nbytes = C.c_size_t()
        native.check(lib.dlwp_spectral_train_workspace(B, Ci, Co, H, W, R, Kc, C.byref(nbytes)), "dlwp_spectral_train_workspace")
        ws = torch.empty(nbytes.value, dtype=torch.uint8, device=x.device)
        y = torch.empty(B * H * W, Co, device=x.device, dtype=torch.float32)
        ri, ro = native.int32_array(rows_in), native.int32_array(rows_out)
        native.check(lib.dlwp_spectral_forward(_p(x), _p(weight), _p(y), B, Ci, Co, H, W, R, Kc, ri, ro, fft_norm, _p(ws), ws.numel(),
                                               _st(x.device)), "dlwp_spectral_forward")
        ctx.save_for_backward(x, weight)
        ctx.geom, ctx.ws = geom, ws
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        B, H, W, rows_in, rows_out, fft_norm = ctx.geom
        Ci, Co, R, Kc = weight.shape[:4]
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw = torch.empty_like(weight)
        ri, ro = native.int32_array(rows_in), native.int32_array(rows_out)
        native.check(native.lib().dlwp_spectral_backward(_p(x), _p(weight), _p(dy), _p(dx), _p(dw), B, Ci, Co, H, W, R, Kc, ri, ro, fft_norm,
                                                         _p(ctx.ws), ctx.ws.numel(), _st(x.device)), "dlwp_spectral_backward")
        return dx, dw, None


def _dense_spectral_weight(wmod, rows: slice, m2: int) -> torch.Tensor:
    """The layer's complex weight ``[C, C, m1, m2]`` as a real view, from the dense ``tensor`` or the Tucker ``core`` + ``factors``."""
    if hasattr(wmod, "tensor"):
        return wmod.tensor[:, :, rows, :m2]
    core = torch.view_as_complex(wmod.core.contiguous())
    ui, uo, ux, uy = (torch.view_as_complex(wmod.factors[k].contiguous()) for k in range(4))
    dense = torch.einsum("abcd,ia,ob,xc,yd->ioxy", core, ui, uo, ux[rows], uy[:m2])
    return torch.view_as_real(dense)


def fno_backbone(module, x: torch.Tensor) -> torch.Tensor:
    """``self.fno(x_t)`` (neuralop FNO.forward: lifting MLP, n_layers x [spectral conv + bias + 1x1 skip, GELU between blocks],
    projection MLP) for ``x [B, Cin, H, W]`` with autograd support."""
    B, Cin, H, W = x.shape
    p = module.fno
    m1, m2 = module._modes_for(H, W)
    rows = centre_crop(module._stored[0], m1)
    rows_in, rows_out = neuralop_row_bins(H, m1)
    geom = (B, H, W, tuple(rows_in), tuple(rows_out), native.FFT_NORM_FORWARD)
    Cn = module.hidden_channels
    a = _Layout.apply(x.to(torch.float32), True, (B, Cin, H, W))
    fcs = p.lifting.fcs
    if len(fcs) == 2:
        a = _Linear.apply(a, fcs[0].weight, fcs[0].bias, None, 1)
        a = _Linear.apply(a, fcs[1].weight, fcs[1].bias, None, 0)
    else:
        a = _Linear.apply(a, fcs[0].weight, fcs[0].bias, None, 0)
    blocks = p.fno_blocks
    for l in range(module.n_layers):
        spec = _Spectral.apply(a, _dense_spectral_weight(blocks.convs.weight[l], rows, m2), geom)
        # x = gelu(spectral(x) + bias_l + skip_l(x)); no activation after the last block
        a = _Linear.apply(a, blocks.fno_skips[l].weight, blocks.convs.bias[l].reshape(Cn), spec, 1 if l < module.n_layers - 1 else 0)
    fcs = p.projection.fcs
    a = _Linear.apply(a, fcs[0].weight, fcs[0].bias, None, 1)
    a = _Linear.apply(a, fcs[1].weight, fcs[1].bias, None, 0)
    return _Layout.apply(a, False, (B, module.prognostic_channels, H, W))


def fno_rollout(module, constants, prescribed, prognostic) -> torch.Tensor:
    """The autoregressive loop of ``fno.py:78-106`` (working upstream form ``fourcastnet.py:502-506``) with autograd."""
    ctx = module.context_size
    outs = []
    for t in range(ctx, prognostic.shape[1]):
        if t == ctx:
            prog_t = prognostic[:, 0:ctx]
        else:
            own = torch.stack(outs, dim=1)[:, -ctx:]
            prog_t = torch.cat([prognostic[:, t - ctx:ctx], own], dim=1) if t - ctx < ctx else own
        parts = []
        if constants is not None:
            parts.append(constants[:, 0])
        if prescribed is not None:
            parts.append(prescribed[:, t - ctx:t].flatten(1, 2))
        parts.append(prog_t.flatten(1, 2))
        outs.append(prog_t[:, -1] + fno_backbone(module, torch.cat(parts, dim=1)))
    return torch.stack(outs, dim=1)
