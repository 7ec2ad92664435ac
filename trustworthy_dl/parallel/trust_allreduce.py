"""Trust-weighted gradient aggregation with fused anomaly statistics (kernels K1-K3).

"Gradient aggregation" is named by the reference (README.md:31; ``optimizer_step`` docstring,
distributed_trainer.py:442) but not implemented; gradient checks are a post-hoc host pass
(attack_detector.py:109-141).  North star: the reduction itself computes, per rank, the bucket L2
norm, the cosine to the coordinate-wise median and the clip factor, weights by current trust,
and reads peer buckets directly over NVLink.

Backends (same Python API, same numbers up to fp rounding):

* ``symm``  — NCCL world on B200s: buckets live in symmetric memory; csrc/trust_allreduce.cu does
  pack+norm (K1), a two-shot reduce where rank r P2P-loads slice r from every peer, runs the
  per-element median network, accumulates dot/|med|^2, and stores the weighted slice to every
  peer (K2); coefficients come from a device kernel fed by the device-resident trust vector (K3).
* ``gather`` — gloo / CPU: all-gather the buckets and run the PyTorch oracle
  (:func:`trustworthy_dl.ops.reference_impl.trust_weighted_reduce`).
* ``local`` — world size 1: statistics only.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.distributed as dist

from ..ops.reference_impl import robust_coefficients, trust_weighted_reduce
from . import comm

__all__ = ["TrustAllReduce"]


class TrustAllReduce:
    def __init__(self, numel: int, dtype: torch.dtype, device: torch.device, clip_ratio: float = 2.0,
                 norm_ratio_max: float = 2.5, backend: Optional[str] = None):
        self.numel, self.dtype, self.device = int(numel), dtype, device
        self.clip_ratio, self.norm_ratio_max = clip_ratio, norm_ratio_max
        self.world, self.rank = comm.world_size(), comm.rank()
        if backend is None:
            if self.world == 1:
                backend = "local"
            elif device.type == "cuda" and dist.get_backend() == "nccl":
                backend = "symm"
            else:
                backend = "gather"
        self.backend = backend
        self._symm = None
        if backend == "symm":
            from .symm import SymmetricTrustReducer
            self._symm = SymmetricTrustReducer(self.numel, dtype, device, clip_ratio, norm_ratio_max)
            self.buffer = self._symm.buffer
        else:
            self.buffer = torch.zeros(self.numel, dtype=dtype, device=device)
        self.last_stats: Optional[torch.Tensor] = None        # [W, 4]: norm, dot, |med|^2, cos
        self.last_coeff: Optional[torch.Tensor] = None

    def reduce_(self, trust_weights: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """In-place: ``buffer <- sum_i c_i * g_i``.  Returns ``(stats[W,4], coeff[W])`` (host tensors
        for gather/local; device tensors for symm)."""
        if self.backend == "symm":
            stats, coeff = self._symm.reduce_(trust_weights)
        elif self.backend == "gather":
            W = self.world
            flat = self.buffer.float()
            gathered = [torch.empty_like(flat) for _ in range(W)]
            dist.all_gather(gathered, flat)
            norms = torch.stack([g.norm() for g in gathered])
            coeff = robust_coefficients(norms, trust_weights, self.clip_ratio, self.norm_ratio_max)
            reduced, stats = trust_weighted_reduce(gathered, coeff, 0.0, norms)
            self.buffer.copy_(reduced.to(self.dtype))
        else:
            g = self.buffer.float()
            norm = g.norm()
            coeff = robust_coefficients(norm.reshape(1), trust_weights[:1], self.clip_ratio, self.norm_ratio_max)
            stats = torch.stack([norm, norm * norm, norm * norm, torch.ones_like(norm)]).reshape(1, 4)
            if float(coeff[0]) != 1.0:
                self.buffer.mul_(float(coeff[0]))
        self.last_stats, self.last_coeff = stats, coeff
        return stats, coeff
