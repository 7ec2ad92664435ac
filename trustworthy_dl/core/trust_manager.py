"""Trust layer: per-node trust scores and the node-health state machine.

Capability parity with the reference trust layer (``/root/reference/trust_manager.py``):
``NodeStatus`` (:18-23), ``TrustScore`` (:25-32), ``NodeMetrics`` (:34-42) and every public
method of ``TrustManager`` (:44-398).  The design is different on purpose:

* State is held in flat numpy vectors (``value / last_updated / count / decay / recovery /
  status / metrics[n, 6]``), not a dict of dataclasses.  The same layout is mirrored on the
  GPU by the ``trust`` / ``status`` vectors of :class:`trustworthy_dl.parallel.symm.SymmetricTrustReducer`:
  the float32 trust vector the fused trust-weighted all-reduce reads and ``finalize_trust_kernel``
  (csrc/trust_allreduce.cu) updates.
* The clock is injectable.  The reference multiplies the old trust by
  ``exp(-decay_rate * wall_clock_dt)`` (:113-114) which makes decisions timing dependent; the
  default clock here is a logical clock that only advances when the trainer says so.
* Both constructor forms are accepted: the code form
  ``TrustManager(num_nodes, trust_threshold, initial_trust, max_history)`` (:49-50) and the
  README form ``TrustManager(initial_trust=, decay_rate=, recovery_rate=)`` (README.md:71-75).
* ``recovery_rate`` (never read by the reference, Q9) speeds up the EMA while a demoted node is
  climbing back; with the default value the trajectory is identical to the reference.
* ``mark_compromised`` records the true previous trust (Q10).
"""
from __future__ import annotations

import json
import logging
import math
from collections import defaultdict, deque
from dataclasses import dataclass
from enum import Enum
from typing import Callable, Dict, Iterable, List, Optional

import numpy as np

logger = logging.getLogger(__name__)

__all__ = [
    "NodeStatus", "TrustScore", "NodeMetrics", "TrustManager", "LogicalClock",
    "STATUS_CODES", "METRIC_FIELDS", "DEFAULT_TRUST_WEIGHTS",
]


class NodeStatus(Enum):
    TRUSTED = "trusted"
    SUSPICIOUS = "suspicious"
    COMPROMISED = "compromised"
    RECOVERING = "recovering"
    OFFLINE = "offline"


#: integer codes shared with the device-side trust kernel (finalize_trust_kernel in csrc/trust_allreduce.cu)
STATUS_CODES: Dict[NodeStatus, int] = {
    NodeStatus.TRUSTED: 0, NodeStatus.SUSPICIOUS: 1, NodeStatus.COMPROMISED: 2,
    NodeStatus.RECOVERING: 3, NodeStatus.OFFLINE: 4,
}
_CODE_TO_STATUS = {v: k for k, v in STATUS_CODES.items()}

METRIC_FIELDS = ("output_deviation", "gradient_consistency", "communication_latency",
                 "resource_utilization", "error_rate", "uptime")
_METRIC_DEFAULTS = np.array([0.0, 1.0, 0.0, 0.0, 0.0, 1.0], dtype=np.float64)
DEFAULT_TRUST_WEIGHTS = {
    "output_deviation": 0.3, "gradient_consistency": 0.3, "communication_latency": 0.1,
    "resource_utilization": 0.1, "error_rate": 0.15, "uptime": 0.05,
}
DEFAULT_DECAY_RATE = 0.01
DEFAULT_RECOVERY_RATE = 0.005


@dataclass
class TrustScore:
    """Snapshot of one node's trust record (reference trust_manager.py:25-32)."""
    value: float
    last_updated: float
    update_count: int
    decay_rate: float = DEFAULT_DECAY_RATE
    recovery_rate: float = DEFAULT_RECOVERY_RATE


@dataclass
class NodeMetrics:
    """Behavioural inputs of the trust formula (reference trust_manager.py:34-42)."""
    output_deviation: float = 0.0
    gradient_consistency: float = 1.0
    communication_latency: float = 0.0
    resource_utilization: float = 0.0
    error_rate: float = 0.0
    uptime: float = 1.0


class LogicalClock:
    """Deterministic clock: time only moves when :meth:`advance` is called."""

    def __init__(self, start: float = 0.0):
        self.now = float(start)

    def advance(self, dt: float) -> None:
        self.now += float(dt)

    def __call__(self) -> float:
        return self.now


class _ScoreView:
    """Dict-like live view so ``tm.trust_scores[i].value`` keeps working."""

    def __init__(self, tm: "TrustManager"):
        self._tm = tm

    def __getitem__(self, node_id: int) -> TrustScore:
        tm = self._tm
        if not tm._known(node_id):
            raise KeyError(node_id)
        return TrustScore(float(tm._value[node_id]), float(tm._last[node_id]),
                          int(tm._count[node_id]), float(tm._decay[node_id]),
                          float(tm._recovery[node_id]))

    def __contains__(self, node_id) -> bool:
        return self._tm._known(node_id)

    def __len__(self) -> int:
        return self._tm.num_nodes

    def __iter__(self):
        return iter(range(self._tm.num_nodes))

    def keys(self):
        return range(self._tm.num_nodes)

    def values(self):
        return [self[i] for i in range(self._tm.num_nodes)]

    def items(self):
        return [(i, self[i]) for i in range(self._tm.num_nodes)]


class _StatusView(_ScoreView):
    def __getitem__(self, node_id: int) -> NodeStatus:
        if not self._tm._known(node_id):
            raise KeyError(node_id)
        return _CODE_TO_STATUS[int(self._tm._status[node_id])]

    def __setitem__(self, node_id: int, status: NodeStatus) -> None:
        self._tm._status[node_id] = STATUS_CODES[status]

    def get(self, node_id, default=None):
        return self[node_id] if self._tm._known(node_id) else default


class _MetricsView(_ScoreView):
    def __getitem__(self, node_id: int) -> NodeMetrics:
        if not self._tm._known(node_id):
            raise KeyError(node_id)
        return NodeMetrics(*[float(x) for x in self._tm._metrics[node_id]])


class TrustManager:
    """Vectorised trust scoring + node status FSM.

    Parameters follow the union of the reference's code and README signatures.
    ``clock`` is any zero-argument callable returning seconds; ``None`` selects a
    :class:`LogicalClock` (reproducible).  Pass ``time.time`` for the reference's wall-clock
    behaviour.
    """

    #: status cut-offs of reference ``_update_node_status`` (:162-181)
    COMPROMISED_BELOW = 0.3
    RECOVERING_ABOVE = 0.8
    RETRUST_ABOVE = 0.9
    COMPROMISED_PENALTY = 0.1

    def __init__(self, num_nodes: Optional[int] = None, trust_threshold: float = 0.7,
                 initial_trust: float = 1.0, max_history: int = 1000, *,
                 decay_rate: float = DEFAULT_DECAY_RATE,
                 recovery_rate: float = DEFAULT_RECOVERY_RATE,
                 alpha: float = 0.1, clock: Optional[Callable[[], float]] = None,
                 trust_weights: Optional[Dict[str, float]] = None):
        self.trust_threshold = float(trust_threshold)
        self.initial_trust = float(initial_trust)
        self.max_history = int(max_history)
        self.default_decay_rate = float(decay_rate)
        self.default_recovery_rate = float(recovery_rate)
        self.alpha = float(alpha)
        self.clock: Callable[[], float] = clock if clock is not None else LogicalClock()
        self.trust_weights = dict(DEFAULT_TRUST_WEIGHTS if trust_weights is None else trust_weights)
        self._w = np.array([self.trust_weights[k] for k in METRIC_FIELDS], dtype=np.float64)

        self.num_nodes = 0
        self._value = np.zeros(0)
        self._last = np.zeros(0)
        self._count = np.zeros(0, dtype=np.int64)
        self._decay = np.zeros(0)
        self._recovery = np.zeros(0)
        self._status = np.zeros(0, dtype=np.int32)
        self._metrics = np.zeros((0, len(METRIC_FIELDS)))

        self.trust_history: Dict[int, deque] = defaultdict(lambda: deque(maxlen=self.max_history))
        self.performance_history: Dict[int, deque] = defaultdict(lambda: deque(maxlen=self.max_history))
        self.attack_history: Dict[int, List] = defaultdict(list)
        self.trust_scores = _ScoreView(self)
        self.node_status = _StatusView(self)
        self.node_metrics = _MetricsView(self)
        if num_nodes:
            self.resize(int(num_nodes))
        logger.info("TrustManager initialized for %d nodes", self.num_nodes)

    # ------------------------------------------------------------------ sizing
    def resize(self, num_nodes: int) -> None:
        """Grow (never shrink) the node table; the README form sizes lazily via the trainer."""
        old = self.num_nodes
        if num_nodes <= old:
            return
        grow = num_nodes - old
        now = self.clock()
        self._value = np.concatenate([self._value, np.full(grow, self.initial_trust)])
        self._last = np.concatenate([self._last, np.full(grow, now)])
        self._count = np.concatenate([self._count, np.zeros(grow, dtype=np.int64)])
        self._decay = np.concatenate([self._decay, np.full(grow, self.default_decay_rate)])
        self._recovery = np.concatenate([self._recovery, np.full(grow, self.default_recovery_rate)])
        self._status = np.concatenate([self._status, np.zeros(grow, dtype=np.int32)])
        self._metrics = np.concatenate([self._metrics, np.tile(_METRIC_DEFAULTS, (grow, 1))])
        self.num_nodes = num_nodes

    def _known(self, node_id) -> bool:
        return isinstance(node_id, (int, np.integer)) and 0 <= node_id < self.num_nodes

    def initialize_node(self, node_id: int) -> None:
        """Reference :82-90 — reset one node to (initial_trust, TRUSTED, default metrics)."""
        if node_id >= self.num_nodes:
            self.resize(node_id + 1)
        self._value[node_id] = self.initial_trust
        self._last[node_id] = self.clock()
        self._count[node_id] = 0
        self._decay[node_id] = self.default_decay_rate
        self._recovery[node_id] = self.default_recovery_rate
        self._status[node_id] = STATUS_CODES[NodeStatus.TRUSTED]
        self._metrics[node_id] = _METRIC_DEFAULTS

    # ----------------------------------------------------------------- scoring
    def instantaneous_score(self, metrics: np.ndarray) -> np.ndarray:
        """Weighted component sum of reference ``_calculate_trust_score`` (:142-160), vectorised
        over the leading axis of ``metrics[..., 6]``."""
        m = np.asarray(metrics, dtype=np.float64)
        comp = np.stack([
            1.0 - np.minimum(1.0, m[..., 0]),
            m[..., 1],
            1.0 - np.minimum(1.0, m[..., 2] / 10.0),
            np.minimum(1.0, m[..., 3]),
            1.0 - np.minimum(1.0, m[..., 4]),
            m[..., 5],
        ], axis=-1)
        return np.clip(comp @ self._w, 0.0, 1.0)

    def _calculate_trust_score(self, node_id: int, metrics: NodeMetrics) -> float:
        vec = np.array([getattr(metrics, f) for f in METRIC_FIELDS], dtype=np.float64)
        return float(self.instantaneous_score(vec))

    def update_trust_score(self, node_id: int, output_deviation: float,
                           gradient_consistency: float, **kwargs) -> float:
        """One EMA step for one node (reference :92-140). Returns the new trust value."""
        if not self._known(node_id):
            self.initialize_node(node_id)
        m = self._metrics[node_id]
        m[0] = output_deviation
        m[1] = gradient_consistency
        for key, val in kwargs.items():
            if key in METRIC_FIELDS:
                m[METRIC_FIELDS.index(key)] = val
        new = float(self.instantaneous_score(m))
        now = self.clock()
        old = float(self._value[node_id])
        decay = math.exp(-float(self._decay[node_id]) * (now - float(self._last[node_id])))
        final = (1.0 - self.alpha) * old * decay + self.alpha * new
        # Q9: recovery_rate only matters once it is raised above the stock default
        gain = float(self._recovery[node_id]) - DEFAULT_RECOVERY_RATE
        if gain > 0.0 and new > old and int(self._status[node_id]) in (
                STATUS_CODES[NodeStatus.COMPROMISED], STATUS_CODES[NodeStatus.RECOVERING]):
            final += gain * (new - old)
        final = min(1.0, max(0.0, final))
        self._value[node_id] = final
        self._last[node_id] = now
        self._count[node_id] += 1
        self._update_node_status(node_id, final)
        self.trust_history[node_id].append({
            "timestamp": now, "trust_score": final,
            "metrics": dict(zip(METRIC_FIELDS, (float(x) for x in m))),
        })
        logger.debug("Node %d trust updated: %.3f", node_id, final)
        return final

    def update_all(self, output_deviation: Iterable[float],
                   gradient_consistency: Iterable[float]) -> np.ndarray:
        """Vector form of :meth:`update_trust_score` for every node in one call."""
        dev = list(output_deviation)
        gc = list(gradient_consistency)
        for i in range(self.num_nodes):
            self.update_trust_score(i, dev[i], gc[i])
        return self._value.copy()

    def next_status_code(self, current: int, trust: float) -> int:
        """Pure transition function of reference ``_update_node_status`` (:162-181), rule order
        preserved (Q11: the RECOVERING>0.9 gate is shadowed by the ``>= threshold`` rule)."""
        if trust < self.COMPROMISED_BELOW:
            return STATUS_CODES[NodeStatus.COMPROMISED]
        if trust < self.trust_threshold:
            return STATUS_CODES[NodeStatus.SUSPICIOUS]
        if current == STATUS_CODES[NodeStatus.COMPROMISED] and trust > self.RECOVERING_ABOVE:
            return STATUS_CODES[NodeStatus.RECOVERING]
        if current == STATUS_CODES[NodeStatus.RECOVERING] and trust > self.RETRUST_ABOVE:
            return STATUS_CODES[NodeStatus.TRUSTED]
        if trust >= self.trust_threshold:
            return STATUS_CODES[NodeStatus.TRUSTED]
        return current

    def _update_node_status(self, node_id: int, trust_score: float) -> None:
        cur = int(self._status[node_id])
        new = self.next_status_code(cur, trust_score)
        if new != cur:
            logger.info("Node %d status changed: %s -> %s", node_id,
                        _CODE_TO_STATUS[cur].value, _CODE_TO_STATUS[new].value)
            self._status[node_id] = new

    # ------------------------------------------------------------ hard actions
    def mark_compromised(self, node_id: int, attack_type: str = "unknown") -> None:
        """Reference :183-196. Trust drops to 0.1 and the node is COMPROMISED immediately."""
        if not self._known(node_id):
            self.initialize_node(node_id)
        previous = float(self._value[node_id])
        self._status[node_id] = STATUS_CODES[NodeStatus.COMPROMISED]
        self._value[node_id] = self.COMPROMISED_PENALTY
        self.attack_history[node_id].append({
            "timestamp": self.clock(), "attack_type": str(attack_type), "previous_trust": previous,
        })
        logger.warning("Node %d marked as compromised: %s", node_id, attack_type)

    def mark_offline(self, node_id: int) -> None:
        """Liveness hook (the reference defines OFFLINE but nothing ever sets it)."""
        if self._known(node_id):
            self._status[node_id] = STATUS_CODES[NodeStatus.OFFLINE]

    def initiate_recovery(self, node_id: int) -> None:
        """Reference :198-206."""
        if self._known(node_id) and int(self._status[node_id]) == STATUS_CODES[NodeStatus.COMPROMISED]:
            self._status[node_id] = STATUS_CODES[NodeStatus.RECOVERING]
            self._recovery[node_id] = 0.02
            logger.info("Recovery initiated for node %d", node_id)

    def reset_node_trust(self, node_id: int) -> None:
        self.initialize_node(node_id)
        logger.info("Trust reset for node %d", node_id)

    # ----------------------------------------------------------------- queries
    def get_trust_score(self, node_id: int) -> float:
        return float(self._value[node_id]) if self._known(node_id) else 0.0

    def get_node_status(self, node_id: int) -> NodeStatus:
        return _CODE_TO_STATUS[int(self._status[node_id])] if self._known(node_id) else NodeStatus.OFFLINE

    def _with_status(self, status: NodeStatus) -> List[int]:
        return [int(i) for i in np.nonzero(self._status == STATUS_CODES[status])[0]]

    def get_trusted_nodes(self) -> List[int]:
        return self._with_status(NodeStatus.TRUSTED)

    def get_suspicious_nodes(self) -> List[int]:
        return self._with_status(NodeStatus.SUSPICIOUS)

    def get_compromised_nodes(self) -> List[int]:
        return self._with_status(NodeStatus.COMPROMISED)

    def can_assign_task(self, node_id: int) -> bool:
        return self.get_node_status(node_id) in (NodeStatus.TRUSTED, NodeStatus.RECOVERING)

    def select_best_nodes(self, num_nodes: int) -> List[int]:
        """Assignable nodes by descending trust (stable on ties, like ``list.sort``)."""
        avail = [(i, self.get_trust_score(i)) for i in range(self.num_nodes) if self.can_assign_task(i)]
        avail.sort(key=lambda t: t[1], reverse=True)
        return [i for i, _ in avail[:num_nodes]]

    def trust_vector(self) -> np.ndarray:
        """float64 copy of all trust values (node order)."""
        return self._value.copy()

    def status_vector(self) -> np.ndarray:
        return self._status.copy()

    def aggregation_weights(self, exclude_compromised: bool = True) -> np.ndarray:
        """Normalised weights the trust-weighted gradient reduction uses: trust values, zeroed
        for COMPROMISED/OFFLINE nodes, summing to one (uniform if everything is zero)."""
        w = self._value.astype(np.float64).copy()
        if exclude_compromised:
            bad = (self._status == STATUS_CODES[NodeStatus.COMPROMISED]) | \
                  (self._status == STATUS_CODES[NodeStatus.OFFLINE])
            w[bad] = 0.0
        s = w.sum()
        return w / s if s > 0 else np.full(self.num_nodes, 1.0 / max(1, self.num_nodes))

    def calculate_system_trust(self) -> float:
        """Trust-weighted mean of trust, sum(t^2)/sum(t) (reference :259-270)."""
        if self.num_nodes == 0:
            return 0.0
        s = float(self._value.sum())
        return float((self._value ** 2).sum() / s) if s > 0 else 0.0

    def get_trust_statistics(self) -> Dict:
        if self.num_nodes == 0:
            return {}
        v = self._value
        return {
            "mean_trust": float(np.mean(v)), "std_trust": float(np.std(v)),
            "min_trust": float(np.min(v)), "max_trust": float(np.max(v)),
            "system_trust": self.calculate_system_trust(),
            "node_status_counts": {s.value: int(np.sum(self._status == c)) for s, c in STATUS_CODES.items()},
            "total_attacks": int(sum(len(a) for a in self.attack_history.values())),
        }

    def get_node_history(self, node_id: int, limit: int = 100) -> List[Dict]:
        if node_id not in self.trust_history:
            return []
        hist = list(self.trust_history[node_id])
        return hist[-limit:] if limit else hist

    def adaptive_threshold_adjustment(self) -> float:
        """Reference :333-348."""
        mean_trust = self.get_trust_statistics().get("mean_trust", 0.7)
        if mean_trust < 0.5:
            self.trust_threshold = max(0.3, mean_trust - 0.1)
        elif mean_trust > 0.9:
            self.trust_threshold = min(0.8, mean_trust - 0.1)
        else:
            self.trust_threshold += 0.01 * (0.7 - self.trust_threshold)
        return self.trust_threshold

    def predict_node_reliability(self, node_id: int, horizon: int = 10) -> float:
        """Linear-trend extrapolation over the last 10 scores (reference :350-368)."""
        if node_id not in self.trust_history or len(self.trust_history[node_id]) < 5:
            return self.get_trust_score(node_id)
        recent = [e["trust_score"] for e in list(self.trust_history[node_id])[-10:]]
        x = np.arange(len(recent))
        slope, intercept = np.polyfit(x, recent, 1)
        return float(np.clip(slope * (len(recent) + horizon) + intercept, 0.0, 1.0))

    def get_recommendations(self) -> List[str]:
        recs: List[str] = []
        stats = self.get_trust_statistics()
        if stats.get("mean_trust", 1.0) < 0.6:
            recs.append("System trust is low - consider investigating compromised nodes")
        if len(self.get_compromised_nodes()) > self.num_nodes * 0.3:
            recs.append("High number of compromised nodes - check security measures")
        if stats.get("total_attacks", 0) > 10:
            recs.append("Frequent attacks detected - strengthen attack detection")
        suspicious = self.get_suspicious_nodes()
        if suspicious:
            recs.append(f"Monitor suspicious nodes: {suspicious}")
        return recs

    # ------------------------------------------------------- persistence / io
    def export_trust_data(self, filepath: str) -> None:
        """JSON export with the reference's schema (:302-331)."""
        data = {
            "trust_scores": {str(i): {"value": float(self._value[i]), "last_updated": float(self._last[i]),
                                      "update_count": int(self._count[i])} for i in range(self.num_nodes)},
            "node_status": {str(i): self.get_node_status(i).value for i in range(self.num_nodes)},
            "trust_history": {str(i): list(h) for i, h in self.trust_history.items()},
            "attack_history": {str(i): a for i, a in self.attack_history.items()},
            "statistics": self.get_trust_statistics(),
        }
        with open(filepath, "w") as f:
            json.dump(data, f, indent=2)
        logger.info("Trust data exported to %s", filepath)

    def state_dict(self) -> Dict:
        """Everything needed to resume with identical decisions."""
        return {
            "num_nodes": self.num_nodes, "trust_threshold": self.trust_threshold,
            "initial_trust": self.initial_trust, "alpha": self.alpha,
            "value": self._value.tolist(), "last": self._last.tolist(), "count": self._count.tolist(),
            "decay": self._decay.tolist(), "recovery": self._recovery.tolist(),
            "status": self._status.tolist(), "metrics": self._metrics.tolist(),
            "clock": float(self.clock()),
            "attack_history": {int(k): list(v) for k, v in self.attack_history.items()},
            "trust_history": {int(k): list(v) for k, v in self.trust_history.items()},
        }

    def load_state_dict(self, sd: Dict) -> None:
        self.num_nodes = 0
        for name in ("_value", "_last", "_decay", "_recovery"):
            setattr(self, name, np.zeros(0))
        self._count = np.zeros(0, dtype=np.int64)
        self._status = np.zeros(0, dtype=np.int32)
        self._metrics = np.zeros((0, len(METRIC_FIELDS)))
        self.trust_threshold = float(sd["trust_threshold"])
        self.initial_trust = float(sd["initial_trust"])
        self.alpha = float(sd.get("alpha", self.alpha))
        if isinstance(self.clock, LogicalClock):
            self.clock.now = float(sd.get("clock", 0.0))
        self.resize(int(sd["num_nodes"]))
        self._value[:] = sd["value"]
        self._last[:] = sd["last"]
        self._count[:] = sd["count"]
        self._decay[:] = sd["decay"]
        self._recovery[:] = sd["recovery"]
        self._status[:] = sd["status"]
        self._metrics[:] = np.asarray(sd["metrics"]).reshape(self.num_nodes, len(METRIC_FIELDS))
        self.attack_history = defaultdict(list, {int(k): list(v) for k, v in sd.get("attack_history", {}).items()})
        self.trust_history = defaultdict(lambda: deque(maxlen=self.max_history))
        for k, v in sd.get("trust_history", {}).items():
            self.trust_history[int(k)].extend(v)

    def cleanup(self) -> None:
        logger.info("TrustManager cleanup completed")
