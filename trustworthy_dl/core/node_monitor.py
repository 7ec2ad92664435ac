"""Running per-node expectations used by the trust inputs.

The reference imports ``NodeMonitor`` (distributed_trainer.py:20,80) and reads
``get_expected_mean/std`` (:234-235) and ``get_expected_gradient_norms`` (:259) but ships no
implementation and nothing ever feeds it.  This one keeps exponentially weighted
running moments of each stage's activation mean/std and of every parameter's gradient norm; the
trainer feeds it after each *clean* step.  Expectations are ``None``/``[]`` until
``warmup`` samples were seen, which reproduces the reference's "not enough history" branches
(:237-239, :261-262).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

__all__ = ["NodeMonitor"]


class NodeMonitor:
    def __init__(self, momentum: float = 0.9, warmup: int = 5):
        self.momentum = float(momentum)
        self.warmup = int(warmup)
        self._mean: Dict[int, float] = {}
        self._std: Dict[int, float] = {}
        self._n_out: Dict[int, int] = {}
        self._gnorms: Dict[int, np.ndarray] = {}
        self._gvar: Dict[int, np.ndarray] = {}
        self._n_grad: Dict[int, int] = {}
        self.band = 2.0                    # expectation = mean - band * std (a conservative floor)
        self.latency: Dict[int, float] = {}

    # ---- feeders -----------------------------------------------------------------
    def record_output(self, node_id: int, mean: float, std: float) -> None:
        n = self._n_out.get(node_id, 0)
        if n == 0:
            self._mean[node_id], self._std[node_id] = float(mean), float(std)
        else:
            m = self.momentum
            self._mean[node_id] = m * self._mean[node_id] + (1 - m) * float(mean)
            self._std[node_id] = m * self._std[node_id] + (1 - m) * float(std)
        self._n_out[node_id] = n + 1

    def record_gradient_norms(self, node_id: int, norms: Sequence[float]) -> None:
        arr = np.asarray(list(norms), dtype=np.float64)
        n = self._n_grad.get(node_id, 0)
        old = self._gnorms.get(node_id)
        if n == 0 or old is None or old.shape != arr.shape:
            self._gnorms[node_id] = arr.copy()
            self._gvar[node_id] = np.zeros_like(arr)
            n = 0
        else:
            m = self.momentum
            self._gvar[node_id] = m * self._gvar[node_id] + (1 - m) * (arr - old) ** 2
            self._gnorms[node_id] = m * old + (1 - m) * arr
        self._n_grad[node_id] = n + 1

    def record_latency(self, node_id: int, seconds: float) -> None:
        self.latency[node_id] = float(seconds)

    # ---- readers (the interface the reference trainer calls) ------------------------
    def get_expected_mean(self, node_id: int) -> Optional[float]:
        return self._mean.get(node_id) if self._n_out.get(node_id, 0) >= self.warmup else None

    def get_expected_std(self, node_id: int) -> Optional[float]:
        if self._n_out.get(node_id, 0) < self.warmup:
            return None
        return max(self._std[node_id], 1e-12)

    def get_expected_gradient_norms(self, node_id: int) -> List[float]:
        """Conservative per-parameter norm floor: running mean minus ``band`` running std (never
        below 10 % of the mean).  The trust formula only penalises norms *below* expectation
        (reference :268), so ordinary step-to-step jitter of a healthy node must stay above it."""
        if self._n_grad.get(node_id, 0) < self.warmup:
            return []
        mean = self._gnorms[node_id]
        floor = np.maximum(mean - self.band * np.sqrt(self._gvar[node_id]), 0.1 * mean)
        return floor.tolist()

    def reset_node(self, node_id: int) -> None:
        for d in (self._mean, self._std, self._n_out, self._gnorms, self._gvar, self._n_grad, self.latency):
            d.pop(node_id, None)

    def state_dict(self) -> Dict:
        return {"mean": dict(self._mean), "std": dict(self._std), "n_out": dict(self._n_out),
                "gnorms": {k: v.tolist() for k, v in self._gnorms.items()},
                "gvar": {k: v.tolist() for k, v in self._gvar.items()}, "n_grad": dict(self._n_grad)}

    def load_state_dict(self, sd: Dict) -> None:
        self._mean = {int(k): float(v) for k, v in sd["mean"].items()}
        self._std = {int(k): float(v) for k, v in sd["std"].items()}
        self._n_out = {int(k): int(v) for k, v in sd["n_out"].items()}
        self._gnorms = {int(k): np.asarray(v, dtype=np.float64) for k, v in sd["gnorms"].items()}
        self._gvar = {int(k): np.asarray(v, dtype=np.float64) for k, v in sd.get("gvar", {}).items()}
        for k, v in self._gnorms.items():
            self._gvar.setdefault(k, np.zeros_like(v))
        self._n_grad = {int(k): int(v) for k, v in sd["n_grad"].items()}
