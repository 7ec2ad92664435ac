"""Trust-driven task reassignment (reference distributed_trainer.py:324-380).

The reference picks ``argmax trust`` among trusted nodes (its comment says "and lowest load" but
load is ignored, Q6), "moves" the partition by aliasing a python object, and estimates migration
time as ``numel * 4 B / 1 GiB/s + 2 s``.  Here the choice is load-aware, stages really change
owner (see :class:`~trustworthy_dl.core.partition.StageMap`), and the time model can use the
measured NVLink peer-copy rate instead of the reference's 1 GiB/s guess."""
from __future__ import annotations

import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

from .trust_manager import TrustManager

__all__ = ["ReassignmentPlanner", "estimate_migration_time", "NVLINK_PEER_BYTES_PER_S"]

NVLINK_PEER_BYTES_PER_S = 770e9           # one-direction peer copy measured with bench/p2p_probe.py on this pool (0.9 TB/s is the link rate)
REFERENCE_BYTES_PER_S = float(1024 ** 3)    # reference assumption (:360)


def estimate_migration_time(numel: int, bytes_per_element: int = 4, model: str = "reference",
                            setup_s: Optional[float] = None) -> float:
    """``model='reference'`` reproduces distributed_trainer.py:354-365 exactly
    (``numel*4/2^30 + 2.0``); ``model='nvlink'`` uses the measured peer bandwidth and a 1 ms setup."""
    if model == "reference":
        return numel * bytes_per_element / REFERENCE_BYTES_PER_S + (2.0 if setup_s is None else setup_s)
    return numel * bytes_per_element / NVLINK_PEER_BYTES_PER_S + (1e-3 if setup_s is None else setup_s)


@dataclass
class ReassignmentPlanner:
    trust_manager: TrustManager
    load_weight: float = 0.25          # how much one "average stage" of extra load costs in trust units
    max_attempts: int = 3              # reference TrainingConfig.max_reassignment_attempts (declared, unused there)
    attempts: Dict[int, int] = field(default_factory=dict)

    def choose_target(self, source: int, loads: Sequence[float]) -> Optional[int]:
        """Highest ``trust - load_weight * relative_load`` among TRUSTED nodes other than source."""
        cands = [n for n in self.trust_manager.get_trusted_nodes() if n != source]
        if not cands:
            return None
        mean_load = (sum(loads) / max(1, sum(1 for l in loads if l > 0))) or 1.0
        return max(cands, key=lambda n: (self.trust_manager.get_trust_score(n)
                                         - self.load_weight * loads[n] / mean_load, -n))

    def allowed(self, source: int) -> bool:
        return self.attempts.get(source, 0) < self.max_attempts

    def note_attempt(self, source: int) -> None:
        self.attempts[source] = self.attempts.get(source, 0) + 1
