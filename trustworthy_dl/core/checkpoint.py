"""Checkpoint save **and** load.  The reference only saves (distributed_trainer.py:448-463): one
``torch.save`` of ``{epoch, global_step, model_partitions, optimizers, trust_scores,
attack_history, reassignment_history}`` into a ``checkpoints/`` directory it never creates (Q7),
and there is no resume path.  The schema here is a superset: scheduler state, node statuses and
full trust state, detector histories, verifier/monitor state, stage map and RNG, so a resumed
run makes the same detection decisions.  Multi-process runs write one shard file
per rank plus rank-0 metadata."""
from __future__ import annotations

import os
from typing import Any, Dict, Optional

import torch

__all__ = ["save_checkpoint", "load_checkpoint", "latest_checkpoint"]


def _path(directory: str, step: int, rank: Optional[int]) -> str:
    name = f"checkpoint_step_{step}.pt" if rank is None else f"checkpoint_step_{step}.rank{rank}.pt"
    return os.path.join(directory, name)


def save_checkpoint(trainer, directory: Optional[str] = None, rank: Optional[int] = None) -> str:
    directory = directory or trainer.config.checkpoint_dir
    os.makedirs(directory, exist_ok=True)
    engine = getattr(getattr(trainer, "_backend_impl", None), "engine", None)
    if engine is not None:                       # fused path: weights live in the engine's flat buffers
        engine.export_to_module(trainer.model_partitions[0])
    impl = getattr(trainer, "_backend_impl", None)
    if hasattr(impl, "export_to_model"):         # fused pipeline stages: same, one engine per hosted stage
        impl.export_to_model()
    ckpt: Dict[str, Any] = {
        # ---- reference schema ----
        "epoch": trainer.current_epoch,
        "global_step": trainer.global_step,
        "model_partitions": {k: v.state_dict() for k, v in trainer.model_partitions.items()},
        "optimizers": {k: v.state_dict() for k, v in trainer.optimizers.items()},
        "trust_scores": {i: trainer.trust_manager.get_trust_score(i) for i in range(trainer.config.num_nodes)},
        "attack_history": list(trainer.attack_history),
        "reassignment_history": list(trainer.reassignment_history),
        # ---- superset ----
        "schedulers": {k: v.state_dict() for k, v in trainer.schedulers.items()},
        "trust_manager": trainer.trust_manager.state_dict(),
        "attack_detector": trainer.attack_detector.state_dict(),
        "gradient_verifier": trainer.gradient_verifier.state_dict(),
        "node_monitor": trainer.node_monitor.state_dict(),
        "stage_map": trainer.stage_map.state_dict() if trainer.stage_map is not None else None,
        "training_state": trainer.training_state.value,
        "epoch_complete": bool(getattr(trainer, "_epoch_complete", False)),
        "rng": {"torch": torch.get_rng_state(),
                "cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None},
        "config": dict(trainer.config.__dict__),
        "format_version": 1,
    }
    if engine is not None:
        ckpt["engine"] = {k: v.cpu() for k, v in engine.state_dict().items()}
    if hasattr(impl, "checkpoint_state"):
        ckpt["strategy"] = impl.checkpoint_state()
    path = _path(directory, trainer.global_step, rank)
    tmp = path + ".tmp"
    torch.save(ckpt, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(trainer, path: str, strict: bool = True, restore_rng: bool = True) -> Dict[str, Any]:
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    trainer.current_epoch = int(ckpt["epoch"])
    trainer.global_step = int(ckpt["global_step"])
    impl = getattr(trainer, "_backend_impl", None)
    # ownership first: a run that reassigned stages before it was checkpointed resumes with THAT map, and a rank must host
    # (module on its device, optimizer / stage engine created) what the map gives it before the states below are loaded
    if ckpt.get("stage_map") and trainer.stage_map is not None:
        trainer.stage_map.load_state_dict(ckpt["stage_map"])
        if hasattr(impl, "sync_ownership"):
            impl.sync_ownership()
    for k, sd in ckpt["model_partitions"].items():
        if k in trainer.model_partitions:
            trainer.model_partitions[k].load_state_dict(sd, strict=strict)
    for k, sd in ckpt.get("optimizers", {}).items():
        if k in trainer.optimizers:
            trainer.optimizers[k].load_state_dict(sd)
    for k, sd in ckpt.get("schedulers", {}).items():
        if k in trainer.schedulers:
            trainer.schedulers[k].load_state_dict(sd)
    if "trust_manager" in ckpt:
        trainer.trust_manager.load_state_dict(ckpt["trust_manager"])
    else:                                               # plain reference-schema file
        for i, v in ckpt.get("trust_scores", {}).items():
            trainer.trust_manager._value[int(i)] = float(v)
    if ckpt.get("attack_detector"):
        trainer.attack_detector.load_state_dict(ckpt["attack_detector"])
    if ckpt.get("gradient_verifier"):
        trainer.gradient_verifier.load_state_dict(ckpt["gradient_verifier"])
    if ckpt.get("node_monitor"):
        trainer.node_monitor.load_state_dict(ckpt["node_monitor"])
    trainer.attack_history = list(ckpt.get("attack_history", []))
    trainer.reassignment_history = list(ckpt.get("reassignment_history", []))
    if "training_state" in ckpt:
        trainer.training_state = type(trainer.training_state)(ckpt["training_state"])
    engine = getattr(getattr(trainer, "_backend_impl", None), "engine", None)
    if engine is not None and ckpt.get("engine"):
        engine.load_state_dict({k: v.to(engine.device) if v.dim() else v for k, v in ckpt["engine"].items()})
    if hasattr(impl, "load_checkpoint_state") and ckpt.get("strategy"):
        impl.load_checkpoint_state(ckpt["strategy"])
    if restore_rng and ckpt.get("rng"):
        torch.set_rng_state(ckpt["rng"]["torch"])
        if ckpt["rng"].get("cuda") is not None and torch.cuda.is_available():
            try:
                torch.cuda.set_rng_state_all(ckpt["rng"]["cuda"])
            except Exception:
                pass
    return ckpt


def latest_checkpoint(directory: str, rank: Optional[int] = None) -> Optional[str]:
    if not os.path.isdir(directory):
        return None
    best, best_step = None, -1
    suffix = ".pt" if rank is None else f".rank{rank}.pt"
    for f in os.listdir(directory):
        if f.startswith("checkpoint_step_") and f.endswith(suffix) and (rank is not None or ".rank" not in f):
            try:
                step = int(f[len("checkpoint_step_"):].split(".")[0])
            except ValueError:
                continue
            if step > best_step:
                best, best_step = os.path.join(directory, f), step
    return best
