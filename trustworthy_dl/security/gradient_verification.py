"""``GradientVerifier`` — imported and called by the reference trainer
(distributed_trainer.py:21,81,199-201: ``verify_gradients(grads, node_id, step) -> bool``) but not
shipped.  Semantics from the north star: per-bucket L2 norm, cosine similarity to the
coordinate-wise median of the replicas, clipping.

Two entry points:

* :meth:`verify_gradients` — single-node view (model-parallel stages have no replica to compare
  against): finiteness + L2 norm against a robust running expectation.
* :meth:`verify_rank_stats` — cross-replica view fed by the fused trust-weighted all-reduce,
  which already produced ``(norm_i, dot(g_i, median), |median|^2)`` per rank inside the
  reduction (csrc/trust_allreduce.cu); nothing is recomputed here.
"""
from __future__ import annotations

import logging
from collections import defaultdict, deque
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

logger = logging.getLogger(__name__)

__all__ = ["GradientVerifier"]


class GradientVerifier:
    def __init__(self, norm_ratio_max: float = 4.0, cosine_min: float = 0.0, warmup: int = 5,
                 history: int = 64, clip_norm: float = 0.0, rank_norm_ratio_max: float = 2.5):
        self.norm_ratio_max = float(norm_ratio_max)
        self.cosine_min = float(cosine_min)
        self.warmup = int(warmup)
        self.clip_norm = float(clip_norm)
        self.rank_norm_ratio_max = float(rank_norm_ratio_max)
        self._norms: Dict[int, deque] = defaultdict(lambda: deque(maxlen=history))
        self.last_norm: Dict[int, float] = {}
        self.rejections: List[Dict] = []

    # ------------------------------------------------------------------ norms
    @staticmethod
    def total_norm(grads: Sequence[torch.Tensor]) -> float:
        """L2 norm over a list of tensors with a single host sync."""
        if not grads:
            return 0.0
        if grads[0].is_cuda:
            from ..ops.multi_tensor import multi_tensor_l2norm
            return float(multi_tensor_l2norm(list(grads)).item())
        sq = torch.stack([g.detach().float().pow(2).sum() for g in grads]).sum()
        return float(sq.sqrt().item())

    def expected_norm(self, node_id: int) -> Optional[float]:
        h = self._norms[node_id]
        return float(np.median(h)) if len(h) >= self.warmup else None

    def verify_norm(self, norm: float, node_id: int, step: int) -> bool:
        """True = valid.  Accepted norms extend the history; rejected ones do not (a poisoner
        cannot drag the expectation along)."""
        self.last_norm[node_id] = norm
        if not np.isfinite(norm):
            self.rejections.append({"node_id": node_id, "step": step, "reason": "non_finite", "norm": norm})
            return False
        exp = self.expected_norm(node_id)
        if exp is not None and exp > 0 and norm > self.norm_ratio_max * exp:
            self.rejections.append({"node_id": node_id, "step": step, "reason": "norm_ratio",
                                    "norm": norm, "expected": exp})
            return False
        self._norms[node_id].append(norm)
        return True

    def verify_gradients(self, gradients: Sequence[torch.Tensor], node_id: int, step: int) -> bool:
        if not gradients:
            return True
        return self.verify_norm(self.total_norm(gradients), node_id, step)

    # ------------------------------------------------------- cross-replica view
    def verify_rank_stats(self, norms: Sequence[float], dots: Sequence[float], med_sq: float,
                          step: int = 0) -> np.ndarray:
        """Per-rank validity from the statistics the fused reduction emits.  A rank is invalid
        when its gradient is non-finite, points away from the coordinate-wise median
        (cosine < ``cosine_min``) or is ``rank_norm_ratio_max`` x larger than the median norm."""
        norms = np.asarray(norms, dtype=np.float64)
        dots = np.asarray(dots, dtype=np.float64)
        W = norms.size
        ok = np.isfinite(norms) & np.isfinite(dots)
        med_norm = np.sqrt(max(float(med_sq), 0.0))
        with np.errstate(divide="ignore", invalid="ignore"):
            cos = np.where((norms > 0) & (med_norm > 0), dots / (norms * med_norm), 1.0)
        if W >= 3:
            ok &= cos >= self.cosine_min
        fin = np.sort(norms[np.isfinite(norms)])
        ref = float(fin[(fin.size - 1) // 2]) if fin.size else 0.0          # lower median (robust for W = 2)
        if W >= 2 and ref > 0:
            ok &= norms <= self.rank_norm_ratio_max * ref
        for r in np.nonzero(~ok)[0]:
            self.rejections.append({"node_id": int(r), "step": step, "reason": "rank_stats",
                                    "norm": float(norms[r]), "cosine": float(cos[r])})
        self.last_cosines = cos
        return ok

    def clip_factor(self, norm: float) -> float:
        if self.clip_norm <= 0 or norm <= self.clip_norm:
            return 1.0
        return self.clip_norm / norm

    def reset_node(self, node_id: int) -> None:
        self._norms.pop(node_id, None)

    def state_dict(self) -> Dict:
        return {"norms": {int(k): list(v) for k, v in self._norms.items()}}

    def load_state_dict(self, sd: Dict) -> None:
        self._norms.clear()
        for k, v in sd.get("norms", {}).items():
            self._norms[int(k)].extend(v)
