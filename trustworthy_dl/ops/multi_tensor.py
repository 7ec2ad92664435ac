"""Multi-tensor reductions: one launch + one sync for the norms of a whole parameter list
(the reference pays one ``.item()`` sync per parameter tensor, distributed_trainer.py:256,312)."""
from __future__ import annotations

from typing import List, Sequence

import torch

from . import _lib

__all__ = ["multi_tensor_l2norm", "per_tensor_l2norms"]


def _table(tensors: Sequence[torch.Tensor]):
    dev = tensors[0].device
    ptrs = torch.tensor([t.data_ptr() for t in tensors], dtype=torch.int64).to(dev, non_blocking=True)
    nums = torch.tensor([t.numel() for t in tensors], dtype=torch.int64).to(dev, non_blocking=True)
    return ptrs, nums


def per_tensor_l2norms(tensors: List[torch.Tensor]) -> torch.Tensor:
    """fp32 vector of per-tensor L2 norms (device tensor for CUDA inputs)."""
    if not tensors:
        return torch.zeros(0)
    if not tensors[0].is_cuda:
        return torch.stack([t.detach().float().norm() for t in tensors])
    ts = [t.detach() if t.is_contiguous() else t.detach().contiguous() for t in tensors]
    dt = ts[0].dtype
    assert all(t.dtype == dt for t in ts), "mixed dtypes: call per dtype group"
    ptrs, nums = _table(ts)
    out = torch.zeros(len(ts), dtype=torch.float32, device=ts[0].device)
    _lib.call("tdl_multi_sqnorm", _lib.ptr(ptrs), _lib.ptr(nums), _lib.i32(len(ts)), _lib.i32(_lib.dtype_code(ts[0])),
              _lib.ptr(out), _lib.stream_ptr())
    return out.sqrt_()


def multi_tensor_l2norm(tensors: List[torch.Tensor]) -> torch.Tensor:
    """Global L2 norm over a list of tensors as a 0-d tensor."""
    if not tensors:
        return torch.zeros(())
    if not tensors[0].is_cuda:
        return torch.stack([t.detach().float().pow(2).sum() for t in tensors]).sum().sqrt()
    groups = {}
    for t in tensors:
        groups.setdefault(t.dtype, []).append(t)
    sq = sum(per_tensor_l2norms(g).pow(2).sum() for g in groups.values())
    return sq.sqrt()
