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

__all__ = ["NodeMonitor", "HeartbeatMonitor"]


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


class HeartbeatMonitor:
    """Host-side liveness view of the ranks (the reference defines ``NodeStatus.OFFLINE`` but has no
    heartbeat, timeout or straggler detection, so nothing ever sets it).

    Every rank reports ``(step, wall_time)`` once per step -- the multi-process strategies already gather a small decision
    vector per step, the heartbeat rides in it -- and every rank runs the same rule on the same gathered data:

    * a rank whose step counter lags the median by more than ``max_step_lag`` steps, or
    * a rank whose report is older than ``timeout_s`` seconds relative to the newest report

    is declared offline (``on_offline`` callback, by default ``TrustManager.mark_offline``); a rank that reports again
    within bounds is returned to service through ``on_online``.  The device-side counterpart is the watchdog of the
    exchange kernel (``csrc/trust_allreduce.cu``), which turns a peer that never arrives into a launch failure.
    """

    def __init__(self, num_nodes: int, max_step_lag: int = 2, timeout_s: float = 60.0, on_offline=None, on_online=None):
        self.num_nodes = int(num_nodes)
        self.max_step_lag = int(max_step_lag)
        self.timeout_s = float(timeout_s)
        self.on_offline = on_offline
        self.on_online = on_online
        self.last_step = np.full(self.num_nodes, -1, dtype=np.int64)
        self.last_time = np.zeros(self.num_nodes, dtype=np.float64)
        self.offline = np.zeros(self.num_nodes, dtype=bool)
        self.events: List[Dict] = []

    def resize(self, num_nodes: int) -> None:
        """Track ``num_nodes`` ranks, keeping what is known about the ones that stay."""
        n, old = int(num_nodes), self.num_nodes
        keep = min(n, old)
        last_step = np.full(n, -1, dtype=np.int64); last_step[:keep] = self.last_step[:keep]
        last_time = np.zeros(n, dtype=np.float64); last_time[:keep] = self.last_time[:keep]
        offline = np.zeros(n, dtype=bool); offline[:keep] = self.offline[:keep]
        self.num_nodes, self.last_step, self.last_time, self.offline = n, last_step, last_time, offline

    @staticmethod
    def beat(step: int, now: float) -> List[float]:
        """The two numbers a rank contributes to the per-step gather."""
        return [float(step), float(now)]

    def observe(self, steps: Sequence[float], times: Sequence[float]) -> List[int]:
        """Feed one gathered heartbeat (index = rank; ``nan`` = no report this round).  Returns the ranks that are
        offline after the update."""
        steps = np.asarray(steps, dtype=np.float64)
        times = np.asarray(times, dtype=np.float64)
        if steps.size != self.num_nodes:
            self.resize(steps.size)          # strategies adopt the launch's world size after construction
        seen = ~np.isnan(steps)
        self.last_step[seen] = steps[seen].astype(np.int64)
        self.last_time[seen] = times[seen]
        known = self.last_step >= 0
        if not known.any():
            return []
        med = float(np.median(self.last_step[known]))
        newest = float(self.last_time[known].max())
        for n in range(self.num_nodes):
            if not known[n]:
                continue
            late = (med - self.last_step[n]) > self.max_step_lag or (newest - self.last_time[n]) > self.timeout_s
            if late and not self.offline[n]:
                self.offline[n] = True
                self.events.append({"node_id": n, "event": "offline", "step": int(med), "lag": float(med - self.last_step[n]),
                                    "silence_s": float(newest - self.last_time[n])})
                if self.on_offline is not None:
                    self.on_offline(n)
            elif not late and self.offline[n]:
                self.offline[n] = False
                self.events.append({"node_id": n, "event": "online", "step": int(med)})
                if self.on_online is not None:
                    self.on_online(n)
        return [int(n) for n in np.nonzero(self.offline)[0]]

    def state_dict(self) -> Dict:
        return {"last_step": self.last_step.tolist(), "last_time": self.last_time.tolist(), "offline": self.offline.tolist(),
                "events": list(self.events)}

    def load_state_dict(self, sd: Dict) -> None:
        self.last_step = np.asarray(sd["last_step"], dtype=np.int64)
        self.last_time = np.asarray(sd["last_time"], dtype=np.float64)
        self.offline = np.asarray(sd["offline"], dtype=bool)
        self.events = list(sd.get("events", []))
