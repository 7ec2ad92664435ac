"""Host-side wrapper of the parameter all-gather with fused integrity checks (K9, csrc/param_gather.cu): pulls every
shard out of the peers' buffers into a local full buffer and, in the same pass, produces per-shard checksums and drift
norms.  The training paths call the kernel directly with device-side verdicts (``GPT2Engine._zero_update`` under
``zero: 1``, ``FusedPipelineStrategy._migrate``); this form returns the checksums to the host (tests, tools)."""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from ..ops import _lib
from ..ops._lib import i32, i64, ptr, stream_ptr

__all__ = ["gather_shards"]


def gather_shards(shard_ptrs: List[int], W: int, dtype: torch.dtype, dst: torch.Tensor, prev: Optional[torch.Tensor],
                  shard_elems: int):
    """Raw kernel call; returns (checksums [W] python ints, drift_sq [W] float64 tensor)."""
    sums = torch.zeros(W, 2, dtype=torch.int64, device=dst.device)
    drift = torch.zeros(W, dtype=torch.float64, device=dst.device)
    arr = (ctypes.c_void_p * W)(*shard_ptrs)
    _lib.call("tdl_param_gather", arr, i32(W), i32(_lib.DTYPE_CODES[dtype]), ptr(dst), ptr(prev), i64(shard_elems),
              ptr(sums), ptr(drift), stream_ptr())
    s = sums.cpu().tolist()
    checks = [((b & 0xFFFFFFFF) << 32) | (a & 0xFFFFFFFF) for a, b in s]
    return checks, drift
