"""Peer-mapped ("symmetric") buffers + the device-side trust-weighted reduction driver.

SURVEY 5.9: every rank allocates identical buffers, exchanges handles once, and kernels then
dereference peer pointers directly over NVLink.  Allocation and the handle exchange go through
``torch.distributed._symmetric_memory`` (cuMem + fabric/posix handles under the hood); everything
on the step path is our own kernels from csrc/trust_allreduce.cu.  With world size 1 the same code
runs on plain local buffers (the "exchange" kernel only waits on its own flag).
"""
from __future__ import annotations

import ctypes
import logging
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import i32, i64, ptr, stream_ptr

logger = logging.getLogger(__name__)

__all__ = ["SymmetricBuffer", "SymmetricTrustReducer", "TRUST_PARAM_DEFAULTS"]

#: alpha, threshold, compromised_below, recovering_above, retrust_above, penalty, clip_ratio,
#: norm_ratio_max, cosine_min, w_dev, w_gc, base_score (= 0.1 latency + 0 util + 0.15 err + 0.05 uptime)
TRUST_PARAM_DEFAULTS = [0.1, 0.7, 0.3, 0.8, 0.9, 0.1, 2.0, 2.5, 0.0, 0.3, 0.3, 0.30]


class SymmetricBuffer:
    """A tensor every rank allocates identically plus the peer pointers to all copies."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, zero: bool = True):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.handle = None
        if self.world == 1:
            self.tensor = torch.zeros(numel, dtype=dtype, device=device)
            self.peer_ptrs = [self.tensor.data_ptr()]
            self.multicast_ptr = 0
            return
        import torch.distributed._symmetric_memory as symm_mem
        self.tensor = symm_mem.empty(numel, dtype=dtype, device=device)
        if zero:
            self.tensor.zero_()
        group = dist.group.WORLD
        try:
            self.handle = symm_mem.rendezvous(self.tensor, group=group)
        except TypeError:
            self.handle = symm_mem.rendezvous(self.tensor, group.group_name)
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.multicast_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        torch.cuda.synchronize()
        dist.barrier()

    def ptr_array(self):
        arr = (ctypes.c_void_p * len(self.peer_ptrs))(*self.peer_ptrs)
        return arr


class SymmetricTrustReducer:
    """Drives K1-K3 for one flat gradient bucket.  All methods only enqueue kernels on the current
    stream (graph-capturable); ``reduce_`` is the convenience path that also returns host copies."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, clip_ratio: float = 2.0,
                 norm_ratio_max: float = 2.5, cosine_min: float = 0.0, trust_threshold: float = 0.7):
        assert device.type == "cuda"
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        W = self.world
        vec = 8 if dtype == torch.bfloat16 else 4
        self.vec = vec
        self.numel = int(numel)
        padded = (numel + vec * W - 1) // (vec * W) * (vec * W)
        self.padded = padded
        self.bucket = SymmetricBuffer(padded, dtype, device)
        self.buffer = self.bucket.tensor[:numel]
        pad_bytes = int(_lib.lib().tdl_exchange_pad_bytes())
        self.pad = SymmetricBuffer((pad_bytes + 7) // 8, torch.int64, device)
        self.dtype = dtype
        self.device = device
        # device-resident state
        self.sumsq = torch.zeros(1, dtype=torch.float64, device=device)
        self.partials = torch.zeros(W + 1, dtype=torch.float64, device=device)
        self.epochs = torch.zeros(2, dtype=torch.int32, device=device)
        self.coeff = torch.zeros(W, dtype=torch.float32, device=device)
        self.norms = torch.zeros(W, dtype=torch.float32, device=device)
        self.stats = torch.zeros(W, 4, dtype=torch.float32, device=device)
        self.flags = torch.zeros(W, dtype=torch.int32, device=device)
        self.trust = torch.ones(W, dtype=torch.float32, device=device)
        self.status = torch.zeros(W, dtype=torch.int32, device=device)
        params = list(TRUST_PARAM_DEFAULTS)
        params[1], params[6], params[7], params[8] = trust_threshold, clip_ratio, norm_ratio_max, cosine_min
        self.params = (ctypes.c_float * 12)(*params)
        self._bucket_ptrs = self.bucket.ptr_array()
        self._pad_ptrs = self.pad.ptr_array()
        nvec = padded // vec
        per = nvec // W
        self.vec_begin, self.vec_end = self.rank * per, (self.rank + 1) * per

    # ---- enqueue-only pieces -------------------------------------------------------------------
    def enqueue_local_norm(self) -> None:
        """sum of squares of the (already packed) local bucket -> self.sumsq (generic path)."""
        from ..ops.stats import tensor_statistics_device  # noqa: F401  (moments kernel reused)
        sq = self.buffer.float().pow(2).sum(dtype=torch.float64)
        self.sumsq.copy_(sq.reshape(1))

    def enqueue_reduce(self, update_trust: bool = True, out_dev: Optional[torch.Tensor] = None) -> None:
        W, r = self.world, self.rank
        _lib.call("tdl_exchange", self._pad_ptrs, i32(r), i32(W), i32(0), ptr(self.sumsq), i32(1), ptr(self.epochs),
                  i32(1), stream_ptr())
        my_pad = ctypes.c_void_p(self.pad.peer_ptrs[r])
        _lib.call("tdl_trust_coefficients", my_pad, i32(W), ptr(self.trust), ptr(self.status), self.params,
                  ptr(self.coeff), ptr(self.norms), stream_ptr())
        _lib.call("tdl_trust_reduce", self._bucket_ptrs, self._bucket_ptrs, i32(W), i32(_lib.DTYPE_CODES[self.dtype]),
                  i64(self.vec_begin), i64(self.vec_end), ptr(self.coeff), ptr(self.partials), stream_ptr())
        _lib.call("tdl_exchange", self._pad_ptrs, i32(r), i32(W), i32(1), ptr(self.partials), i32(W + 1),
                  ptr(self.epochs), i32(1), stream_ptr())
        _lib.call("tdl_trust_finalize", my_pad, i32(W), ptr(self.norms), ptr(out_dev), self.params, ptr(self.stats),
                  ptr(self.trust), ptr(self.status), ptr(self.flags), i32(1 if update_trust else 0), stream_ptr())

    # ---- convenience (eager DP strategy) --------------------------------------------------------
    def load_trust(self, trust_values, status_codes) -> None:
        self.trust.copy_(torch.as_tensor(trust_values, dtype=torch.float32))
        self.status.copy_(torch.as_tensor(status_codes, dtype=torch.int32))

    def reduce_(self, trust_weights: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """Eager path: weights given by the host trust manager (already zero for compromised ranks)."""
        self.trust.copy_(trust_weights.to(self.device, torch.float32) * self.world)
        self.status.zero_()
        self.enqueue_local_norm()
        self.enqueue_reduce(update_trust=False)
        return self.stats.clone(), self.coeff.clone()
