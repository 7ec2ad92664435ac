"""Activation / gradient statistics (kernel K5): the detector's 12 statistics of
reference attack_detector.py:185-200 without moving the tensor to the host.

CUDA path (csrc/stats.cu): one streaming pass accumulates shifted power sums, min/max, L1, Linf;
exact median / p25 / p75 come from an 8-bit-digit radix select over order-preserving integer
keys (2 passes for bf16/fp16, 4 for fp32).  Result: 12 float64 on the device; only those cross
PCIe when the host detector wants them.
"""
from __future__ import annotations

import torch

from . import _lib
from .reference_impl import NUM_STATS, STAT_NAMES, tensor_statistics as _ref_stats, moments_raw as _ref_moments

__all__ = ["tensor_statistics", "tensor_statistics_device", "mean_std", "STAT_NAMES", "NUM_STATS"]


def tensor_statistics_device(x: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
    """12 statistics as a float64 *device* tensor (no host sync)."""
    assert x.is_cuda
    x = x.detach()
    if not x.is_contiguous():
        x = x.contiguous()
    n = x.numel()
    if out is None:
        out = torch.empty(NUM_STATS, dtype=torch.float64, device=x.device)
    nbytes = int(_lib.lib().tdl_stats_scratch_bytes(_lib.i64(n)))
    scr = _lib.scratch(x.device, nbytes, "stats")
    _lib.call("tdl_tensor_stats", _lib.ptr(x), _lib.i32(_lib.dtype_code(x)), _lib.i64(n), _lib.ptr(out),
              _lib.ptr(scr), ctypes_size(scr.numel()), _lib.stream_ptr())
    return out


def ctypes_size(v):
    import ctypes
    return ctypes.c_size_t(int(v))


def tensor_statistics(x: torch.Tensor) -> torch.Tensor:
    """12 statistics in ``STAT_NAMES`` order (float64 tensor; device tensor for CUDA inputs)."""
    if x.is_cuda:
        return tensor_statistics_device(x)
    return _ref_stats(x)


def mean_std(x: torch.Tensor, unbiased: bool = True):
    """(mean, std) as python floats with one sync — the trust inputs of reference
    distributed_trainer.py:242-243 (torch's default unbiased std)."""
    if x.is_cuda:
        s = tensor_statistics_device(x)[:2].cpu()
        mean, std = float(s[0]), float(s[1])
        n = x.numel()
        if unbiased and n > 1:
            std *= (n / (n - 1)) ** 0.5
        return mean, std
    xf = x.detach().float()
    return float(xf.mean()), float(xf.std(unbiased=unbiased))
