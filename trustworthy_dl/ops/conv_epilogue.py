"""K13 wrapper: per-channel affine (+residual) + ReLU as one kernel, with autograd.

``fused_bias_relu(x, bias)`` replaces conv-bias + ReLU of the VGG stages; ``fused_affine_add_relu`` is the
BN-folded / residual form used by ResNet blocks in eval or frozen-BN mode.  CPU tensors use plain PyTorch."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._lib import i32, i64, ptr, stream_ptr


def _layout(x: torch.Tensor):
    N, C = x.shape[0], x.shape[1]
    HW = x.numel() // (N * C)
    if x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last) and not x.is_contiguous():
        return N, C, HW, 1
    return N, C, HW, 0


class _AffineAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, shift, residual, relu):
        if not x.is_cuda:
            shape = (1, -1) + (1,) * (x.dim() - 2)
            y = x.float()
            if scale is not None:
                y = y * scale.view(shape)
            if shift is not None:
                y = y + shift.view(shape)
            if residual is not None:
                y = y + residual.float()
            y = (torch.relu(y) if relu else y).to(x.dtype)
        else:
            N, C, HW, cl = _layout(x)
            if not cl:
                x = x.contiguous()
                residual = residual.contiguous() if residual is not None else None
            y = torch.empty_like(x)
            _lib.call("tdl_conv_epilogue_fwd", ptr(x), ptr(scale), ptr(shift), ptr(residual), ptr(y), i64(N), i64(C), i64(HW),
                      i32(_lib.dtype_code(x)), i32(cl), i32(relu), stream_ptr())
        ctx.save_for_backward(x, y, scale)
        ctx.relu, ctx.has_res, ctx.has_shift = relu, residual is not None, shift is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, scale = ctx.saved_tensors
        C = x.shape[1]
        if not x.is_cuda:
            g = dy.float() * ((y > 0).float() if ctx.relu else 1.0)
            shape = (1, -1) + (1,) * (x.dim() - 2)
            dims = [d for d in range(x.dim()) if d != 1]
            dx = (g * scale.view(shape) if scale is not None else g).to(x.dtype)
            dscale = (g * x.float()).sum(dims) if scale is not None else None
            dshift = g.sum(dims) if ctx.has_shift else None
            return dx, dscale, dshift, (g.to(x.dtype) if ctx.has_res else None), None
        N, C, HW, cl = _layout(x)
        dy = dy.contiguous(memory_format=torch.channels_last) if cl else dy.contiguous()
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if ctx.has_res else None
        dscale = torch.zeros(C, dtype=torch.float32, device=x.device) if scale is not None else None
        dshift = torch.zeros(C, dtype=torch.float32, device=x.device) if ctx.has_shift else None
        _lib.call("tdl_conv_epilogue_bwd", ptr(dy), ptr(y), ptr(x if scale is not None else None), ptr(scale), ptr(dx), ptr(dres),
                  ptr(dscale), ptr(dshift), i64(N), i64(C), i64(HW), i32(_lib.dtype_code(x)), i32(cl), i32(ctx.relu), stream_ptr())
        return dx, dscale, dshift, dres, None


def fused_affine_add_relu(x: torch.Tensor, scale: Optional[torch.Tensor], shift: Optional[torch.Tensor],
                          residual: Optional[torch.Tensor] = None, relu: bool = True) -> torch.Tensor:
    """relu(x * scale[c] + shift[c] + residual); scale/shift are fp32 per-channel vectors."""
    return _AffineAct.apply(x, scale, shift, residual, relu)


def fused_bias_relu(x: torch.Tensor, bias: torch.Tensor) -> torch.Tensor:
    return _AffineAct.apply(x, None, bias.float(), None, True)
