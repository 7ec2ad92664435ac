"""YAML <-> dataclasses.  The reference documents a YAML schema (README.md:111-132:
``model{name,size}``, ``training{batch_size,learning_rate,num_epochs}``,
``distributed{num_nodes,parallelism}``, ``security{trust_threshold,attack_detection,
gradient_verification}``) and parses ``--config`` but never reads it (experiment_runner.py:604,
Q17).  This loader implements the documented schema plus optional ``attack``, ``experiment`` and
``trust`` sections."""
from __future__ import annotations

from dataclasses import asdict, fields
from typing import Any, Dict, Optional, Tuple

import yaml

from .core.distributed_trainer import TrainingConfig

__all__ = ["load_yaml", "training_config_from_dict", "load_training_config", "dump_training_config",
           "experiment_config_from_dict"]

_PARALLELISMS = ("model", "pipeline", "data", "tensor")


def load_yaml(path: str) -> Dict[str, Any]:
    with open(path) as f:
        data = yaml.safe_load(f) or {}
    if not isinstance(data, dict):
        raise ValueError(f"{path}: top level must be a mapping")
    return data


def training_config_from_dict(d: Dict[str, Any], **overrides) -> TrainingConfig:
    model = d.get("model", {}) or {}
    tr = d.get("training", {}) or {}
    ds = d.get("distributed", {}) or {}
    sec = d.get("security", {}) or {}
    par = ds.get("parallelism", "model")
    if par not in _PARALLELISMS:
        raise ValueError(f"distributed.parallelism must be one of {_PARALLELISMS}, got {par!r}")
    kw: Dict[str, Any] = dict(
        model_name=model.get("name", "gpt2"), model_size=model.get("size"),
        model_kwargs=dict(model.get("kwargs", {}) or {}),
        dataset_name=tr.get("dataset", d.get("dataset", {}).get("name", "openwebtext") if isinstance(d.get("dataset"), dict) else "openwebtext"),
        batch_size=int(tr.get("batch_size", 32)), learning_rate=float(tr.get("learning_rate", 5e-5)),
        num_epochs=int(tr.get("num_epochs", 10)), num_nodes=int(ds.get("num_nodes", 4)), parallelism=par,
        trust_threshold=float(sec.get("trust_threshold", 0.7)),
        attack_detection_enabled=bool(sec.get("attack_detection", True)),
        gradient_verification_enabled=bool(sec.get("gradient_verification", True)),
        parameter_monitoring_enabled=bool(sec.get("parameter_monitoring", False)),
    )
    opt = {"checkpoint_interval": tr, "steps_per_epoch": tr, "seq_len": tr, "optimizer": tr, "weight_decay": tr,
           "dtype": tr, "clip_norm": tr, "micro_batches": ds, "tp_size": ds, "pp_size": ds, "dp_size": ds,
           "max_reassignment_attempts": sec, "device": ds, "lr_schedule": tr, "checkpoint_dir": tr, "seed": tr,
           "engine": tr, "migration_model": ds, "zero": ds,
           "detector_require_evidence": sec, "heartbeat_interval": ds, "pp_virtual": ds, "probation": sec, "detector_min_history": sec}
    for name, section in opt.items():
        if name in section:
            kw[name] = section[name]
    kw.update(overrides)
    return TrainingConfig.from_kwargs(**kw)


def load_training_config(path: str, **overrides) -> TrainingConfig:
    return training_config_from_dict(load_yaml(path), **overrides)


def dump_training_config(cfg: TrainingConfig) -> str:
    d = {"model": {"name": cfg.model_name, "size": cfg.model_size},
         "training": {"batch_size": cfg.batch_size, "learning_rate": cfg.learning_rate, "num_epochs": cfg.num_epochs,
                      "dataset": cfg.dataset_name},
         "distributed": {"num_nodes": cfg.num_nodes, "parallelism": cfg.parallelism},
         "security": {"trust_threshold": cfg.trust_threshold, "attack_detection": cfg.attack_detection_enabled,
                      "gradient_verification": cfg.gradient_verification_enabled}}
    return yaml.safe_dump(d, sort_keys=False)


def experiment_config_from_dict(d: Dict[str, Any], **overrides):
    """Build an :class:`ExperimentConfig` from the same YAML (optional ``experiment``/``attack`` sections)."""
    from .experiments.runner import ExperimentConfig
    tc = training_config_from_dict(d)
    ex = d.get("experiment", {}) or {}
    at = d.get("attack", {}) or {}
    kw = dict(experiment_name=ex.get("name", f"{tc.model_name}_{tc.dataset_name}_nodes{tc.num_nodes}"),
              model_name=tc.model_name, dataset_name=tc.dataset_name, num_nodes=tc.num_nodes,
              num_epochs=tc.num_epochs, batch_size=tc.batch_size, learning_rate=tc.learning_rate,
              attack_enabled=bool(at.get("enabled", False)), attack_start_epoch=int(at.get("start_epoch", 2)),
              attack_intensity=float(at.get("intensity", 0.5)), trust_threshold=tc.trust_threshold,
              save_interval=int(ex.get("save_interval", 100)), output_dir=ex.get("output_dir", "results"))
    for name in ("attack_types", "target_nodes", "gradient_mode"):
        if name in at:
            kw[name] = at[name]
    for name in ("steps_per_epoch",):
        if name in ex:
            kw[name] = ex[name]
    kw.update(overrides)
    cfg = ExperimentConfig(**kw)
    cfg.training_overrides = {k: getattr(tc, k) for k in ("model_size", "model_kwargs", "parallelism", "seq_len",
                                                          "dtype", "optimizer", "device", "micro_batches",
                                                          "parameter_monitoring_enabled")}
    return cfg
