"""Peer-mapped ("symmetric") buffers + the device-side trust-weighted reduction driver.

SURVEY 5.9: every rank allocates identical buffers, exchanges handles once, and kernels then
dereference peer pointers directly over NVLink.  Allocation and the handle exchange go through
``torch.distributed._symmetric_memory`` (cuMem + fabric/posix handles under the hood); everything
on the step path is our own kernels from csrc/trust_allreduce.cu.  With world size 1 the same code
runs on plain local buffers (the "exchange" kernel only waits on its own flag).

A reducer may span the whole world (data parallelism) or a subset of ranks (the replicas of one
pipeline stage): the arena is always allocated world-symmetric, the kernels are simply handed the
peer pointers of the participating ranks, so group membership can change at run time (stage
reassignment) without a new rendezvous.
"""
from __future__ import annotations

import ctypes
import logging
import os
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import f32, i32, i64, ptr, stream_ptr

logger = logging.getLogger(__name__)

__all__ = ["SymmetricBuffer", "SymmetricTrustReducer", "TRUST_PARAM_DEFAULTS"]

#: alpha, threshold, compromised_below, recovering_above, retrust_above, penalty, clip_ratio,
#: norm_ratio_max, cosine_min, w_dev, w_gc, base_score (= 0.1 latency + 0 util + 0.15 err + 0.05 uptime)
TRUST_PARAM_DEFAULTS = [0.1, 0.7, 0.3, 0.8, 0.9, 0.1, 2.0, 2.5, 0.0, 0.3, 0.3, 0.30]


class SymmetricBuffer:
    """A tensor every rank allocates identically plus the peer pointers to all copies."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, zero: bool = True, local: bool = False):
        """``local``: plain device memory even inside a multi-rank job (a group of one has no peers to map; the
        symmetric rendezvous is a world collective and must not be entered by one rank alone)."""
        self.world = dist.get_world_size() if dist.is_initialized() and not local else 1
        self.rank = dist.get_rank() if dist.is_initialized() and not local else 0
        self.handle = None
        self.itemsize = torch.empty(0, dtype=dtype).element_size()
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

    def ptr_array(self, ranks: Optional[Sequence[int]] = None, offset_elems: int = 0):
        """``void*[len(ranks)]`` of the copies held by ``ranks`` (default: every rank), advanced by ``offset_elems``."""
        ranks = range(len(self.peer_ptrs)) if ranks is None else ranks
        off = offset_elems * self.itemsize
        return (ctypes.c_void_p * len(ranks))(*[self.peer_ptrs[r] + off for r in ranks])


class SymmetricTrustReducer:
    """Drives K1-K4 for one flat gradient bucket (or a sequence of sub-buckets of it).  All ``enqueue_*``
    methods only launch kernels on the current stream (graph-capturable); ``reduce_`` is the convenience
    path that also returns host copies.

    ``ranks``: global ranks taking part (default: the world).  ``arena`` / ``arena_offset`` / ``pad`` /
    ``pad_index``: place the bucket inside an existing world-symmetric allocation (pipeline stages share one
    arena per rank, stage ``s`` always at the same offset, so the replicas of a stage find each other wherever
    the stage currently runs)."""

    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, clip_ratio: float = 2.0,
                 norm_ratio_max: float = 2.5, cosine_min: float = 0.0, trust_threshold: float = 0.7,
                 ranks: Optional[Sequence[int]] = None, arena: Optional[SymmetricBuffer] = None, arena_offset: int = 0,
                 pad: Optional[SymmetricBuffer] = None, pad_index: int = 0):
        assert device.type == "cuda"
        gworld = dist.get_world_size() if dist.is_initialized() else 1
        grank = dist.get_rank() if dist.is_initialized() else 0
        self.ranks = list(range(gworld)) if ranks is None else list(ranks)
        assert grank in self.ranks, "this rank is not a member of the reduction group"
        self.world = len(self.ranks)
        self.rank = self.ranks.index(grank)
        W = self.world
        vec = 8 if dtype == torch.bfloat16 else 4
        self.vec = vec
        self.numel = int(numel)
        padded = (numel + vec * W - 1) // (vec * W) * (vec * W)
        self.padded = padded
        self.dtype = dtype
        self.device = device
        if arena is None:
            self.bucket = SymmetricBuffer(padded, dtype, device, local=W == 1)
            self.arena_offset = 0
        else:
            assert arena_offset % vec == 0 and arena_offset + padded <= arena.tensor.numel()
            self.bucket = arena
            self.arena_offset = int(arena_offset)
        self.full = self.bucket.tensor[self.arena_offset:self.arena_offset + padded]
        self.buffer = self.full[:numel]
        self.pad_bytes = int(_lib.lib().tdl_exchange_pad_bytes())
        self.pad_words = (self.pad_bytes + 7) // 8
        if pad is None:
            self.pad = SymmetricBuffer(self.pad_words, torch.int64, device, local=W == 1)
            self.pad_index = 0
        else:
            self.pad, self.pad_index = pad, int(pad_index)
        # device-resident state
        self.sumsq = torch.zeros(1, dtype=torch.float64, device=device)
        self.partials = torch.zeros(W + 1, dtype=torch.float64, device=device)
        self.epochs = torch.zeros(2, dtype=torch.int32, device=device)
        self.coeff = torch.full((W,), 1.0 / W, dtype=torch.float32, device=device)
        self.norms = torch.zeros(W, dtype=torch.float32, device=device)
        self.stats = torch.zeros(W, 4, dtype=torch.float32, device=device)
        self.flags = torch.zeros(W, dtype=torch.int32, device=device)
        self.trust = torch.ones(W, dtype=torch.float32, device=device)
        self.status = torch.zeros(W, dtype=torch.int32, device=device)
        self.acc_sq = torch.zeros(W, dtype=torch.float64, device=device)        # whole-step squared norms (bucketed mode)
        self.acc_bad = torch.zeros(W, dtype=torch.int32, device=device)         # rejected in some bucket this step
        self._nothing = torch.zeros(1, dtype=torch.float64, device=device)
        params = list(TRUST_PARAM_DEFAULTS)
        params[1], params[6], params[7], params[8] = trust_threshold, clip_ratio, norm_ratio_max, cosine_min
        self.params = (ctypes.c_float * 12)(*params)
        # NVLS (multimem) paths exist only for whole-world groups: the multicast object spans every rank of the allocation.
        # K4 (plain mean all-reduce) goes through the switch whenever it can (TDL_K4_MULTIMEM=0: P2P).  K2's write-back
        # keeps W peer stores by default: the reduction already RECEIVES (W-1)/W of the bucket as load responses, and a
        # multicast store loops the own slice back through the switch on top of that (RX (2W-1)/W vs 2(W-1)/W of the
        # bucket), which measured slower at W = 2; TDL_TR_MULTIMEM=1 selects the single multimem.st.
        has_mc = self.bucket.multicast_ptr != 0 and W == gworld and W > 1
        self.use_multimem = has_mc and os.environ.get("TDL_TR_MULTIMEM", "0") == "1"
        self.k4_multimem = has_mc and os.environ.get("TDL_K4_MULTIMEM", "1") != "0"
        self.set_group(self.ranks)

    # ---- group membership ------------------------------------------------------------------------
    def set_group(self, ranks: Sequence[int]) -> None:
        """(Re)bind the peer pointers to ``ranks`` (same length as at construction)."""
        assert len(ranks) == self.world
        grank = dist.get_rank() if dist.is_initialized() else 0
        self.ranks = list(ranks)
        self.rank = self.ranks.index(grank)
        idx = self.ranks if self.bucket.world > 1 else [0]
        self._bucket_ptrs = self.bucket.ptr_array(idx, self.arena_offset)
        self._pad_ptrs = self.pad.ptr_array(idx, self.pad_index * self.pad_words)
        self._my_pad = ctypes.c_void_p(self.pad.peer_ptrs[idx[self.rank]] + self.pad_index * self.pad_words * 8)
        mc = ctypes.c_void_p(self.bucket.multicast_ptr + self.arena_offset * self.bucket.itemsize) if self.bucket.multicast_ptr else None
        self._mc = mc if self.use_multimem else None
        self._mc_k4 = mc if self.k4_multimem else None
        nvec = self.padded // self.vec
        per = nvec // self.world
        self.vec_begin, self.vec_end = self.rank * per, (self.rank + 1) * per

    def slice_of(self, begin: int, numel: int) -> Tuple[int, int]:
        """This rank's share (in 16-byte vectors, relative to the bucket start) of elements [begin, begin+numel)."""
        assert begin % self.vec == 0 and numel % self.vec == 0
        v0, nv = begin // self.vec, numel // self.vec
        per = (nv + self.world - 1) // self.world
        lo = min(self.rank * per, nv)
        hi = min(lo + per, nv)
        return v0 + lo, v0 + hi

    # ---- enqueue-only pieces -------------------------------------------------------------------
    def barrier(self, which: int = 1) -> None:
        _lib.call("tdl_exchange", self._pad_ptrs, i32(self.rank), i32(self.world), i32(which), ptr(self._nothing), i32(0),
                  ptr(self.epochs), i32(0), stream_ptr())

    def enqueue_local_norm(self) -> None:
        """sum of squares of the (already packed) local bucket -> self.sumsq (K1's norm half, via the checked path)."""
        _lib.call("tdl_sumsq", ptr(self.buffer), i32(_lib.DTYPE_CODES[self.dtype]), i64(self.numel), ptr(self.sumsq), stream_ptr())

    def enqueue_bucket(self, begin: int, numel: int, sumsq: Optional[torch.Tensor] = None, overlapped: bool = False,
                       owner_only: bool = False, accumulate: bool = True) -> None:
        """K2 on elements [begin, begin+numel) of the bucket with coefficients from THIS sub-bucket's norms
        (gathered from ``sumsq``, a device double): exchange norms -> coefficients -> reduce -> barrier.
        ``overlapped``: kept for callers; the reduction always uses the 256-thread kernel.
        ``owner_only``: every rank keeps only its own slice of the result (ZeRO-1).
        Statistics accumulate over the calls of a step; :meth:`enqueue_finalize` closes the step."""
        W, r = self.world, self.rank
        sumsq = self.sumsq if sumsq is None else sumsq
        _lib.call("tdl_exchange", self._pad_ptrs, i32(r), i32(W), i32(0), ptr(sumsq), i32(1), ptr(self.epochs), i32(1),
                  stream_ptr())
        _lib.call("tdl_trust_coefficients_acc", self._my_pad, i32(W), ptr(self.trust), ptr(self.status), self.params,
                  ptr(self.coeff), ptr(self.norms), ptr(self.acc_sq) if accumulate else None,
                  ptr(self.acc_bad) if accumulate else None, stream_ptr())
        vb, ve = self.slice_of(begin, numel)
        _lib.call("tdl_trust_reduce_ex", self._bucket_ptrs, self._bucket_ptrs, i32(W), i32(_lib.DTYPE_CODES[self.dtype]),
                  i64(vb), i64(ve), ptr(self.coeff), ptr(self.partials), i32(r if owner_only else -1),
                  None if owner_only else self._mc, i32(256), i32(0), stream_ptr())
        self.barrier(1)                                               # every peer's slice has landed in my bucket

    def enqueue_finalize(self, update_trust: bool = True, out_dev: Optional[torch.Tensor] = None,
                         accumulated: bool = True) -> None:
        """Gather the per-rank dot/|median|^2 partials, compute cosines and flags, run the trust EMA + state machine."""
        W, r = self.world, self.rank
        _lib.call("tdl_exchange", self._pad_ptrs, i32(r), i32(W), i32(1), ptr(self.partials), i32(W + 1),
                  ptr(self.epochs), i32(1), stream_ptr())
        _lib.call("tdl_trust_finalize_acc", self._my_pad, i32(W), ptr(self.norms), ptr(out_dev), self.params, ptr(self.stats),
                  ptr(self.trust), ptr(self.status), ptr(self.flags), i32(1 if update_trust else 0),
                  ptr(self.acc_sq) if accumulated else None, ptr(self.acc_bad) if accumulated else None, stream_ptr())

    def enqueue_reduce(self, update_trust: bool = True, out_dev: Optional[torch.Tensor] = None) -> None:
        """One-bucket form (the whole buffer, norm already in ``self.sumsq``)."""
        self.enqueue_bucket(0, self.padded, accumulate=False)
        self.enqueue_finalize(update_trust, out_dev, accumulated=False)

    def enqueue_sum_allreduce(self, begin: int = 0, numel: Optional[int] = None, scale: Optional[float] = None) -> None:
        """K4: plain mean all-reduce of [begin, begin+numel) (no statistics, no trust weighting) -- through the switch
        (multimem.ld_reduce / multimem.st) when the bucket has a multicast address, over P2P otherwise."""
        W, r = self.world, self.rank
        numel = self.padded if numel is None else numel
        vb, ve = self.slice_of(begin, numel)
        self.barrier(0)                                               # every rank's bucket is packed
        _lib.call("tdl_sum_allreduce", self._bucket_ptrs, self._mc_k4, i32(W), i32(_lib.DTYPE_CODES[self.dtype]), i64(vb), i64(ve),
                  f32(1.0 / W if scale is None else scale), i32(0), stream_ptr())
        self.barrier(1)

    # ---- convenience (eager DP strategy) --------------------------------------------------------
    def reduce_(self, trust_weights: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """Eager path: weights given by the host trust manager (already zero for compromised ranks).  Returns device
        tensors ``(stats[W,4], coeff[W])`` that stay valid until the next call (no clone, no sync)."""
        self.trust.copy_(trust_weights.to(self.device, torch.float32) * self.world, non_blocking=True)
        self.status.zero_()
        self.enqueue_local_norm()
        self.enqueue_reduce(update_trust=False)
        return self.stats, self.coeff
