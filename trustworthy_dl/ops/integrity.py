"""Parameter-integrity primitives (kernel K9 building blocks): 64-bit shard checksum and
``sum((theta - theta_prev)^2)`` with the snapshot refreshed in the same pass."""
from __future__ import annotations

import torch

from . import _lib
from .reference_impl import fletcher64

__all__ = ["checksum64", "drift_sq_and_update"]


def checksum64(t: torch.Tensor) -> int:
    t = t.detach()
    if not t.is_contiguous():
        t = t.contiguous()
    if not t.is_cuda:
        return fletcher64(t)
    nbytes = t.numel() * t.element_size()
    assert nbytes % 4 == 0, "shards are padded to 32-bit words"
    out = torch.zeros(2, dtype=torch.int64, device=t.device)
    _lib.call("tdl_checksum64", _lib.ptr(t), _lib.i64(nbytes // 4), _lib.ptr(out), _lib.stream_ptr())
    s1, s2 = (int(v) & 0xFFFFFFFF for v in out.cpu().tolist())
    return (s2 << 32) | s1


def drift_sq_and_update(cur: torch.Tensor, prev: torch.Tensor) -> float:
    """Returns sum((cur-prev)^2) and overwrites ``prev`` with ``cur``."""
    if not cur.is_cuda:
        d = (cur.float() - prev.float()).pow(2).sum().item()
        prev.copy_(cur)
        return float(d)
    out = torch.zeros(1, dtype=torch.float32, device=cur.device)
    _lib.call("tdl_drift_update", _lib.ptr(cur), _lib.ptr(prev), _lib.i64(cur.numel()), _lib.i32(_lib.dtype_code(cur)),
              _lib.ptr(out), _lib.stream_ptr())
    return float(out.item())
