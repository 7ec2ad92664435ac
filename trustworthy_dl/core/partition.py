"""Stage partitioning.  The reference splits GPT blocks with ``len(h) // num_nodes`` (dropping the
remainder, leaving embeddings/ln_f/head on no stage) and has no ResNet/VGG split at all
(distributed_trainer.py:116-146).  Here every model exposes ``stage_units()`` — an
ordered list of tensor->tensor modules — and the partitioner cuts that list into contiguous,
cost-balanced stages (minimising the heaviest stage; exact DP, the lists are short)."""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import torch.nn as nn

__all__ = ["unit_cost", "balanced_partition", "build_stages", "StageMap"]


def unit_cost(unit: nn.Module, floor: float = 1e4) -> float:
    """Parameter count with a floor so parameter-free units (pools) still count as work."""
    return float(sum(p.numel() for p in unit.parameters())) + floor


def balanced_partition(costs: Sequence[float], num_stages: int) -> List[int]:
    """Boundaries ``b[0]=0 < b[1] < ... < b[S]=len(costs)`` of the contiguous split that minimises
    the maximum stage cost.  Every stage gets at least one unit when ``len(costs) >= num_stages``."""
    n = len(costs)
    S = max(1, min(num_stages, n))
    prefix = [0.0]
    for c in costs:
        prefix.append(prefix[-1] + c)
    INF = float("inf")
    best = [[INF] * (n + 1) for _ in range(S + 1)]
    cut = [[0] * (n + 1) for _ in range(S + 1)]
    best[0][0] = 0.0
    for s in range(1, S + 1):
        for i in range(s, n - (S - s) + 1):
            for j in range(s - 1, i):
                v = max(best[s - 1][j], prefix[i] - prefix[j])
                if v < best[s][i]:
                    best[s][i], cut[s][i] = v, j
    bounds = [n]
    i = n
    for s in range(S, 0, -1):
        i = cut[s][i]
        bounds.append(i)
    bounds.reverse()
    return bounds


def build_stages(model: nn.Module, num_stages: int, cost_fn: Callable[[nn.Module], float] = unit_cost
                 ) -> List[nn.Sequential]:
    if hasattr(model, "stage_units"):
        units = list(model.stage_units())
    elif isinstance(model, nn.Sequential):
        units = list(model.children())            # a user's own sequential model: its children are the units
    else:
        raise TypeError(f"{type(model).__name__} does not expose stage_units() and is not an nn.Sequential; cannot partition")
    bounds = balanced_partition([cost_fn(u) for u in units], num_stages)
    return [nn.Sequential(*units[a:b]) for a, b in zip(bounds[:-1], bounds[1:])]


class StageMap:
    """Which node runs which stage.  Stages keep their order; ownership moves on reassignment."""

    def __init__(self, num_stages: int, num_nodes: int):
        self.num_stages, self.num_nodes = num_stages, num_nodes
        self.owner: List[int] = [s % num_nodes for s in range(num_stages)] if num_stages > num_nodes \
            else list(range(num_stages))

    def stages_of(self, node: int) -> List[int]:
        return [s for s, o in enumerate(self.owner) if o == node]

    def move_all(self, src: int, dst: int) -> List[int]:
        moved = self.stages_of(src)
        for s in moved:
            self.owner[s] = dst
        return moved

    def load(self, node: int, stage_costs: Sequence[float]) -> float:
        return float(sum(stage_costs[s] for s in self.stages_of(node)))

    def state_dict(self) -> Dict:
        return {"owner": list(self.owner), "num_nodes": self.num_nodes}

    def load_state_dict(self, sd: Dict) -> None:
        self.owner = list(sd["owner"])
