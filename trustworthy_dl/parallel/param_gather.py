"""ZeRO-1 style sharded parameters with the integrity checks fused into the all-gather (K9).

Each rank owns shard ``rank`` of the flat bf16 parameter buffer (it ran the optimizer on that shard).
``gather()`` pulls every shard out of the peers' symmetric buffers into the local full buffer and, in the
same pass, produces per-shard checksums and drift norms.  Owners publish the checksum of what they wrote
(``publish()``); a mismatch on any receiver flags ``MODEL_POISONING`` for the owner.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import i32, i64, ptr, stream_ptr
from .symm import SymmetricBuffer

__all__ = ["ShardedParameterGather", "gather_shards"]


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


class ShardedParameterGather:
    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, track_drift: bool = True):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        W = self.world
        self.shard_elems = (numel + W * 8 - 1) // (W * 8) * 8
        self.dtype, self.device = dtype, device
        self.shard = SymmetricBuffer(self.shard_elems, dtype, device)        # what this rank owns / publishes
        self.full = torch.zeros(W * self.shard_elems, dtype=dtype, device=device)
        self.prev = torch.zeros_like(self.full) if track_drift else None
        self.published = torch.zeros(W, dtype=torch.int64, device=device)
        self.numel = numel

    def publish(self) -> int:
        """Owner side: checksum of the local shard, shared with every rank."""
        from ..ops.integrity import checksum64
        cs = checksum64(self.shard.tensor)
        if self.world > 1:
            t = torch.tensor([cs - (1 << 64) if cs >= (1 << 63) else cs], dtype=torch.int64, device=self.device)
            out = torch.empty(self.world, dtype=torch.int64, device=self.device)
            dist.all_gather_into_tensor(out, t)
            self.published = out
        else:
            self.published = torch.tensor([cs - (1 << 64) if cs >= (1 << 63) else cs], dtype=torch.int64, device=self.device)
        return cs

    def gather(self) -> Dict:
        if self.world > 1:
            dist.barrier()
        checks, drift = gather_shards(self.shard.peer_ptrs, self.world, self.dtype, self.full, self.prev, self.shard_elems)
        pub = [(int(v) + (1 << 64)) % (1 << 64) for v in self.published.cpu().tolist()]
        bad = [i for i, (a, b) in enumerate(zip(checks, pub)) if a != b]
        return {"checksums": checks, "published": pub, "mismatched_shards": bad, "drift": drift.sqrt().cpu().tolist()}
