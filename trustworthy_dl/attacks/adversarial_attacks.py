"""Fault injector.  The reference's runner drives an ``AdversarialAttacker`` it does not ship
(experiment_runner.py:23,90-97,157-160,187-188,231,285,597-598; README.md:61): constructor
``AttackConfig(attack_types, target_nodes, intensity, start_step)`` and methods
``activate_attacks / is_active / apply_attacks(batch, batch_idx) / get_attack_statistics /
get_final_statistics / cleanup``.  North star adds gradient and parameter poisoning.

Attack surfaces (all deterministic given ``seed``):

=====================  ==========================================================
``data_poisoning``     label flips + input noise on a fraction ``intensity`` of the batch
``gradient_poisoning`` in-place on a target node's gradients: ``scale`` (x(1+20·i)),
                       ``sign_flip`` (x -(1+4·i)), ``noise`` (+N(0, (4·i·rms)^2))
``model_poisoning``    in-place parameter perturbation, relative size ``intensity``
``output_poisoning``   scales/offsets a stage's activation (a compromised stage)
``backdoor``           stamps a trigger patch / token and forces a target label
=====================  ==========================================================

On CUDA tensors the gradient/parameter mutations are single fused kernels
(csrc/attack_kernels.cu, counter-based RNG) so injecting faults does not perturb timing.
"""
from __future__ import annotations

import logging
from collections import defaultdict
from dataclasses import dataclass, field
from typing import Any, Dict, Iterable, List, Optional, Sequence

import torch

logger = logging.getLogger(__name__)

__all__ = ["AttackConfig", "AdversarialAttacker", "GRADIENT_MODES"]

GRADIENT_MODES = ("scale", "sign_flip", "noise")


@dataclass
class AttackConfig:
    attack_types: List[str] = field(default_factory=lambda: ["gradient_poisoning"])
    target_nodes: List[int] = field(default_factory=lambda: [1])
    intensity: float = 0.5
    start_step: int = 0
    end_step: Optional[int] = None
    gradient_mode: str = "scale"
    frequency: int = 1              # attack every k-th eligible step
    backdoor_target: int = 0
    seed: int = 1234


class AdversarialAttacker:
    def __init__(self, config: Optional[AttackConfig] = None, **kwargs):
        self.config = config if config is not None else AttackConfig(**kwargs)
        if self.config.gradient_mode not in GRADIENT_MODES:
            raise ValueError(f"gradient_mode must be one of {GRADIENT_MODES}")
        self._active = False
        self._gen = torch.Generator().manual_seed(self.config.seed)
        self.counts: Dict[str, int] = defaultdict(int)
        self.per_node: Dict[int, int] = defaultdict(int)
        self.log: List[Dict[str, Any]] = []
        self.samples_poisoned = 0

    # ---------------------------------------------------------------- lifecycle
    def activate_attacks(self) -> None:
        if not self._active:
            logger.warning("adversarial attacks activated: %s on nodes %s", self.config.attack_types,
                           self.config.target_nodes)
        self._active = True

    def deactivate_attacks(self) -> None:
        self._active = False

    def is_active(self) -> bool:
        return self._active

    def _armed(self, kind: str, step: int, node_id: Optional[int] = None) -> bool:
        c = self.config
        if not self._active or kind not in c.attack_types or step < c.start_step:
            return False
        if c.end_step is not None and step >= c.end_step:
            return False
        if node_id is not None and node_id not in c.target_nodes:
            return False
        return (step - c.start_step) % max(1, c.frequency) == 0

    def _record(self, kind: str, step: int, node_id: Optional[int] = None, **extra) -> None:
        self.counts[kind] += 1
        if node_id is not None:
            self.per_node[node_id] += 1
        self.log.append({"type": kind, "step": step, "node_id": node_id, **extra})

    NODE_ATTACKS = ("gradient_poisoning", "model_poisoning", "output_poisoning")

    def injects(self, step: int, node_id: int) -> bool:
        """Would this configuration inject a node-level fault at ``node_id`` in ``step`` (if the node hosts work)?  Unlike
        :meth:`was_attacked` this does not read the log, so every process of a multi-process run -- each has its own
        injector, and only the target's own process logs the injection -- gives the same answer."""
        return any(self._armed(k, step, node_id) for k in self.NODE_ATTACKS)

    def was_attacked(self, step: int, node_id: Optional[int] = None) -> bool:
        return any(e["step"] == step and (node_id is None or e["node_id"] == node_id) for e in self.log)

    # -------------------------------------------------------------- data attacks
    def apply_attacks(self, batch: Dict[str, torch.Tensor], batch_idx: int) -> Dict[str, torch.Tensor]:
        """Data-level attacks on a ``{'input','target'}`` batch (runner call site :187-188).
        With data poisoning disabled (or the attacker idle) the batch is returned untouched."""
        step = batch_idx
        out = batch
        if self._armed("data_poisoning", step):
            out = self._poison_data(dict(out), step)
        if self._armed("backdoor", step):
            out = self._stamp_backdoor(dict(out), step)
        return out

    def _poison_data(self, batch, step):
        x, y = batch["input"], batch["target"]
        n = x.shape[0]
        k = max(1, int(round(self.config.intensity * n)))
        idx = torch.randperm(n, generator=self._gen)[:k].to(x.device)
        if x.is_floating_point():
            x = x.clone()
            noise = torch.randn(x[idx].shape, generator=self._gen).to(x.device, x.dtype)
            x[idx] = x[idx] + self.config.intensity * x.float().std().to(x.dtype) * noise
            y = y.clone()
            ncls = int(y.max().item()) + 1
            y[idx] = (y[idx] + 1 + torch.randint(0, max(1, ncls - 1), (k,), generator=self._gen).to(y.device)) % max(2, ncls)
        else:                                   # token streams: corrupt inputs, shuffle targets
            x = x.clone()
            vocab = int(x.max().item()) + 1
            mask = torch.rand(x[idx].shape, generator=self._gen).to(x.device) < self.config.intensity
            rnd = torch.randint(0, vocab, x[idx].shape, generator=self._gen).to(x.device)
            x[idx] = torch.where(mask, rnd, x[idx])
            y = y.clone()
            y[idx] = y[idx].flip(-1)
        batch["input"], batch["target"] = x, y
        self.samples_poisoned += k
        self._record("data_poisoning", step, None, samples=k)
        return batch

    def _stamp_backdoor(self, batch, step):
        x, y = batch["input"].clone(), batch["target"].clone()
        n = x.shape[0]
        k = max(1, int(round(self.config.intensity * n * 0.5)))
        if x.is_floating_point() and x.dim() == 4:
            x[:k, :, -3:, -3:] = x.max()
            y[:k] = self.config.backdoor_target
        else:
            x[:k, -1] = self.config.backdoor_target
        batch["input"], batch["target"] = x, y
        self._record("backdoor", step, None, samples=k)
        return batch

    # ---------------------------------------------------------- gradient attacks
    def poison_gradients(self, node_id: int, gradients: Iterable[torch.Tensor], step: int) -> bool:
        """In-place poison of one node's gradients; returns True when something was changed."""
        if not self._armed("gradient_poisoning", step, node_id):
            return False
        i, mode = self.config.intensity, self.config.gradient_mode
        seed = self.config.seed + 7919 * step + node_id
        for j, g in enumerate(gradients):
            if g is None:
                continue
            if g.is_cuda:
                from ..ops.attack_ops import poison_inplace
                poison_inplace(g, mode, i, seed + j)
            elif mode == "scale":
                g.mul_(1.0 + 20.0 * i)
            elif mode == "sign_flip":
                g.mul_(-(1.0 + 4.0 * i))
            else:
                rms = g.float().pow(2).mean().sqrt()
                gen = torch.Generator().manual_seed(seed + j)
                g.add_((4.0 * i * rms * torch.randn(g.shape, generator=gen)).to(g.dtype))
        self._record("gradient_poisoning", step, node_id, mode=mode)
        return True

    def poison_parameters(self, node_id: int, params: Iterable[torch.Tensor], step: int) -> bool:
        if not self._armed("model_poisoning", step, node_id):
            return False
        seed = self.config.seed + 104729 * step + node_id
        with torch.no_grad():
            for j, p in enumerate(params):
                if p.is_cuda:
                    from ..ops.attack_ops import poison_inplace
                    poison_inplace(p.data, "noise", self.config.intensity / 4.0, seed + j)
                else:
                    gen = torch.Generator().manual_seed(seed + j)
                    rms = p.float().pow(2).mean().sqrt().clamp_min(1e-3)
                    p.add_((self.config.intensity * rms * torch.randn(p.shape, generator=gen)).to(p.dtype))
        self._record("model_poisoning", step, node_id)
        return True

    def poison_output(self, node_id: int, output: torch.Tensor, step: int) -> torch.Tensor:
        """A compromised stage emitting corrupted activations."""
        if not self._armed("output_poisoning", step, node_id):
            return output
        self._record("output_poisoning", step, node_id)
        return output * (1.0 + 10.0 * self.config.intensity)

    # ------------------------------------------------------------------- stats
    def get_attack_statistics(self) -> Dict[str, Any]:
        return {"active": self._active, "total_attacks": int(sum(self.counts.values())),
                "by_type": dict(self.counts), "by_node": {int(k): v for k, v in self.per_node.items()},
                "samples_poisoned": self.samples_poisoned, "intensity": self.config.intensity,
                "target_nodes": list(self.config.target_nodes)}

    def get_final_statistics(self) -> Dict[str, Any]:
        s = self.get_attack_statistics()
        s["first_attack_step"] = self.log[0]["step"] if self.log else None
        s["last_attack_step"] = self.log[-1]["step"] if self.log else None
        s["attack_types"] = list(self.config.attack_types)
        return s

    def cleanup(self) -> None:
        self._active = False
