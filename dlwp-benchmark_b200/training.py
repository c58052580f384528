"""Training-mode forward / backward of ``FNO2DModule`` / ``TFNO2DModule``: ``scripts/train.py:263-271`` calls
``model(...)``, a loss and ``.backward()``.  With grad mode on, the modules route here instead of the inference plan:

* every FLOP of the backbone runs in ``libdlwp_b200.so`` (fp32 kernel set): ``dlwp_linear_forward/_backward`` (neuralop MLP
  1x1 convolutions, FNO skip convolution + bias + GELU epilogue) and ``dlwp_spectral_forward/_backward`` (dense Fourier layer
  and its adjoint + per-mode weight gradient), wrapped in ``torch.autograd.Function`` objects on channels-last tensors;
* torch autograd only glues: the step-input concatenation and residual frame of the rollout (``fno.py:49-62,78-106``) and,
  for TFNO, the reconstruction of the dense spectral weight from the Tucker core and factors (a 86 k-entry einsum whose
  gradient autograd carries back to ``core`` / ``factors``).

FourCastNet (``AFNONet``, ``fourcastnet.py:214-361``) trains through the same primitives plus ``dlwp_layernorm_forward/_backward``
and ``dlwp_afno_filter_forward/_backward`` (the AFNO2D token mixer: truncated ortho rfft2, block-diagonal complex MLP with ReLU and
softshrink, irfft2, both skips; the backward pass recomputes the spectrum and the hidden layer).  With ``filter: FNO2D`` the token
mixer is the one-layer FNO of ``fourcastnet.py:130-153`` on ``_Linear`` / ``_Spectral``.  Torch glue there: the patch gather /
un-patch permutations and the broadcast of ``pos_embed`` over the batch.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import native
from .plan import afno_kept_region, centre_crop, neuralop_row_bins


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


class _LayerNorm(torch.autograd.Function):
    """``nn.LayerNorm(C, eps)`` over the last axis of ``x [P, C]``."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        x = x.contiguous()
        P, Cn = x.shape
        y = torch.empty_like(x)
        native.check(native.lib().dlwp_layernorm_forward(_p(x), _p(gamma.contiguous()), _p(beta.contiguous()), _p(y), P, Cn, float(eps),
                                                         _st(x.device)), "dlwp_layernorm_forward")
        ctx.save_for_backward(x, gamma)
        ctx.eps = float(eps)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma = ctx.saved_tensors
        dy = dy.contiguous()
        P, Cn = x.shape
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dg, db = torch.empty_like(gamma), torch.empty_like(gamma)
        native.check(native.lib().dlwp_layernorm_backward(_p(dy), _p(x), _p(gamma.contiguous()), _p(dx), _p(dg), _p(db), P, Cn, ctx.eps,
                                                          _st(x.device)), "dlwp_layernorm_backward")
        return dx, dg, db, None


class _AfnoFilter(torch.autograd.Function):
    """``AFNO2D.forward(x) + extra`` (``fourcastnet.py:78-127``; ``extra`` = the Block's second skip) on ``x [B*h*w, C]``."""

    @staticmethod
    def forward(ctx, x, extra, w1, b1, w2, b2, geom):
        B, h, w, rows, cols, nb, lam = geom
        x = x.contiguous()
        Cn = x.shape[1]
        f = w1.shape[3] // w1.shape[2]
        lib = native.lib()
        nbytes = C.c_size_t()
        native.check(lib.dlwp_afno_filter_workspace(B, Cn, h, w, len(rows), cols, f, C.byref(nbytes)), "dlwp_afno_filter_workspace")
        ws = torch.empty(nbytes.value, dtype=torch.uint8, device=x.device)
        y = torch.empty_like(x)
        par = [t.contiguous() for t in (w1, b1, w2, b2)]
        native.check(lib.dlwp_afno_filter_forward(_p(x), _p(extra.contiguous() if extra is not None else None), *(_p(t) for t in par), _p(y),
                                                  B, Cn, h, w, len(rows), cols, native.int32_array(rows), nb, f, float(lam), _p(ws),
                                                  ws.numel(), _st(x.device)), "dlwp_afno_filter_forward")
        ctx.save_for_backward(x, *par)
        ctx.geom, ctx.ws, ctx.f, ctx.has_extra = geom, ws, f, extra is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w1, b1, w2, b2 = ctx.saved_tensors
        B, h, w, rows, cols, nb, lam = ctx.geom
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        grads = [torch.empty_like(t) for t in (w1, b1, w2, b2)]
        native.check(native.lib().dlwp_afno_filter_backward(_p(x), _p(w1), _p(b1), _p(w2), _p(b2), _p(dy), _p(dx), *(_p(g) for g in grads), B,
                                                            x.shape[1], h, w, len(rows), cols, native.int32_array(rows), nb, ctx.f,
                                                            float(lam), _p(ctx.ws), ctx.ws.numel(), _st(x.device)),
                     "dlwp_afno_filter_backward")
        return (dx, dy if ctx.has_extra else None, *grads, None)


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


def _fno2d_filter(module, blk, u: torch.Tensor, t: torch.Tensor, B: int) -> torch.Tensor:
    """``FNO2D.forward(u) + t`` (``fourcastnet.py:145-153``): lifting MLP, one FNO block without activation, projection MLP, ``+ u``."""
    p = blk.filter.fno
    E = module.embed_dim
    m1, m2 = module._fno2d_modes()
    rows_in, rows_out = neuralop_row_bins(module.h, m1)
    geom = (B, module.h, module.w, tuple(rows_in), tuple(rows_out), native.FFT_NORM_FORWARD)
    a = _Linear.apply(u, p.lifting.fcs[0].weight, p.lifting.fcs[0].bias, None, 1)
    a = _Linear.apply(a, p.lifting.fcs[1].weight, p.lifting.fcs[1].bias, None, 0)
    n1 = p.fno_blocks.convs.weight[0].tensor.shape[2]
    spec = _Spectral.apply(a, p.fno_blocks.convs.weight[0].tensor[:, :, centre_crop(n1, m1), :m2], geom)
    a = _Linear.apply(a, p.fno_blocks.fno_skips[0].weight, p.fno_blocks.convs.bias[0].reshape(E), spec, 0)
    a = _Linear.apply(a, p.projection.fcs[0].weight, p.projection.fcs[0].bias, None, 1)
    return _Linear.apply(a, p.projection.fcs[1].weight, p.projection.fcs[1].bias, u + t, 0)


def afno_backbone(module, x: torch.Tensor) -> torch.Tensor:
    """``AFNONet.forward_features`` + head + un-patch (``fourcastnet.py:283-293,345-357``) for ``x [B, Cin, H, W]`` with autograd."""
    B, Cin, H, W = x.shape
    if (H, W) != module.img_size:
        raise AssertionError(f"Input image size ({H}*{W}) doesn't match model ({module.img_size[0]}*{module.img_size[1]}).")
    p1, p2 = module.patch_size
    h, w, E, Cg = module.h, module.w, module.embed_dim, module.out_chans
    # patch gather (PatchEmbed's strided convolution as a Linear on [Cin, p1, p2] patches, :537-542)
    tok = x.to(torch.float32).reshape(B, Cin, h, p1, w, p2).permute(0, 2, 4, 1, 3, 5).reshape(B * h * w, Cin * p1 * p2)
    pos = module.pos_embed.expand(B, h * w, E).reshape(B * h * w, E) if module.use_pos_embed else None
    t = _Linear.apply(tok, module.patch_embed.proj.weight.reshape(E, Cin * p1 * p2), module.patch_embed.proj.bias, pos, 0)
    if module.filter_kind == "AFNO2D":
        rows, cols = afno_kept_region(h, w // 2 + 1, module.hard_thresholding_fraction)
        if not rows or cols <= 0:
            raise NotImplementedError("hard_thresholding_fraction keeps no Fourier mode")
        geom = (B, h, w, tuple(rows), cols, module.num_blocks, module.sparsity_threshold)
    for blk in module.blocks:
        u = _LayerNorm.apply(t, blk.norm1.weight, blk.norm1.bias, blk.norm1.eps)
        if module.filter_kind == "AFNO2D":      # x = filter(norm1(x)) + x with the filter's own "+ input" (:127, :185-187)
            f = blk.filter
            t2 = _AfnoFilter.apply(u, t, f.w1, f.b1, f.w2, f.b2, geom)
        else:
            t2 = _fno2d_filter(module, blk, u, t, B)
        u = _LayerNorm.apply(t2, blk.norm2.weight, blk.norm2.bias, blk.norm2.eps)
        a = _Linear.apply(u, blk.mlp.fc1.weight, blk.mlp.fc1.bias, None, 1)
        t = _Linear.apply(a, blk.mlp.fc2.weight, blk.mlp.fc2.bias, t2, 0)      # x = mlp(norm2(x)) + x
    y = _Linear.apply(t, module.head.weight, None, None, 0)      # no bias; self.norm is never applied (:263)
    return y.reshape(B, h, w, p1, p2, Cg).permute(0, 5, 1, 3, 2, 4).reshape(B, Cg, H, W)


def backbone(module, x: torch.Tensor) -> torch.Tensor:
    return afno_backbone(module, x) if hasattr(module, "patch_embed") else fno_backbone(module, x)


def rollout(module, constants, prescribed, prognostic) -> torch.Tensor:
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
        outs.append(prog_t[:, -1] + backbone(module, torch.cat(parts, dim=1)))
    return torch.stack(outs, dim=1)


fno_rollout = rollout
