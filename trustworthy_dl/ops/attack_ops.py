"""Fault-injection kernels (K16): in-place scale / sign-flip / additive Gaussian noise with a
counter-based RNG, so poisoning a bucket is one launch."""
from __future__ import annotations

import torch

from . import _lib

_MODES = {"scale": 0, "sign_flip": 1, "noise": 2}


def poison_inplace(t: torch.Tensor, mode: str, intensity: float, seed: int) -> None:
    assert t.is_cuda and t.is_contiguous()
    if mode == "scale":
        a, b = 1.0 + 20.0 * intensity, 0.0
    elif mode == "sign_flip":
        a, b = -(1.0 + 4.0 * intensity), 0.0
    else:
        rms = float(t.float().pow(2).mean().sqrt().item())
        a, b = 1.0, 4.0 * intensity * rms
    _lib.call("tdl_poison_inplace", _lib.ptr(t), _lib.i64(t.numel()), _lib.i32(_lib.dtype_code(t)),
              _lib.f32(a), _lib.f32(b), _lib.i64(seed), _lib.stream_ptr())
