"""Attack sweep: the same experiment repeated over attack types and intensities, tabulating what README.md:146-149 asks
for -- detection rate, false positives / missed nodes and detection latency "across various attack intensities" -- from
measured runs (the reference has no such driver; its runner simulates the loss, experiment_runner.py:201-216).

    python -m trustworthy_dl.experiments.sweep --model resnet32 --dataset cifar10 --nodes 2 \\
        --types gradient_poisoning output_poisoning --intensities 0.05 0.1 0.25 0.5 1.0
"""
from __future__ import annotations

import argparse
import json
import logging
from dataclasses import replace
from pathlib import Path
from typing import Any, Dict, List, Optional, Sequence

from .runner import ExperimentConfig, ExperimentRunner

logger = logging.getLogger(__name__)

__all__ = ["attack_sweep", "main"]


def attack_sweep(base: ExperimentConfig, attack_types: Sequence[str], intensities: Sequence[float],
                 gradient_modes: Sequence[str] = ("scale",)) -> List[Dict[str, Any]]:
    """One :class:`ExperimentRunner` run per (attack type, gradient mode, intensity) plus one clean run (false-positive
    reference).  Returns the rows and writes ``sweep_results.json`` / ``sweep_table.md`` under ``base.output_dir``."""
    rows: List[Dict[str, Any]] = []

    def one(name: str, **changes) -> Dict[str, Any]:
        cfg = replace(base, experiment_name=name, **changes)
        res = ExperimentRunner(cfg).run_experiment()
        s = res.get("experiment_summary", {})
        so = s.get("security_outcomes", {}) or {}
        return {"experiment": name, "final_loss": s.get("final_loss"), "attacks_detected": s.get("total_attacks_detected"),
                "reassignments": s.get("reassignments"), "detection_rate": so.get("detection_rate"),
                "mean_detection_latency_steps": so.get("mean_detection_latency_steps"),
                "missed_nodes": so.get("missed_nodes", []), "false_positive_nodes": so.get("false_positive_nodes", []),
                "peak_loss_after_attack": so.get("peak_loss_after_attack"), "recovery_steps": so.get("recovery_steps")}

    clean = one(f"{base.experiment_name}_clean", attack_enabled=False)
    rows.append({"attack_type": "none", "gradient_mode": None, "intensity": 0.0, **clean})
    for kind in attack_types:
        for mode in (gradient_modes if kind == "gradient_poisoning" else (None,)):
            for x in intensities:
                tag = f"{base.experiment_name}_{kind}{'_' + mode if mode else ''}_{x:g}"
                extra = {}
                if kind == "model_poisoning":           # injected (and caught) by the parameter monitor's hooks
                    extra["training_overrides"] = {**base.training_overrides, "parameter_monitoring_enabled": True}
                r = one(tag, attack_enabled=True, attack_types=[kind], attack_intensity=float(x),
                        gradient_mode=mode or base.gradient_mode, **extra)
                rows.append({"attack_type": kind, "gradient_mode": mode, "intensity": float(x), **r})
    out = Path(base.output_dir)
    out.mkdir(parents=True, exist_ok=True)
    (out / "sweep_results.json").write_text(json.dumps(rows, indent=2, default=str))
    head = ["attack", "mode", "intensity", "detection rate", "latency (steps)", "missed", "false positives", "peak loss", "recovery (steps)", "final loss"]

    def fmt(v, spec=".3f"):
        return format(v, spec) if isinstance(v, (int, float)) and not isinstance(v, bool) else ("-" if v in (None, [], "") else str(v))

    lines = ["| " + " | ".join(head) + " |", "|" + "---|" * len(head)]
    for r in rows:
        lines.append("| " + " | ".join([r["attack_type"], r["gradient_mode"] or "-", fmt(r["intensity"], "g"), fmt(r["detection_rate"], ".2f"),
                                        fmt(r["mean_detection_latency_steps"], ".1f"), fmt(r["missed_nodes"]), fmt(r["false_positive_nodes"]),
                                        fmt(r["peak_loss_after_attack"]), fmt(r["recovery_steps"], "d"), fmt(r["final_loss"])]) + " |")
    (out / "sweep_table.md").write_text("\n".join(lines) + "\n")
    return rows


def main(argv: Optional[Sequence[str]] = None) -> int:
    p = argparse.ArgumentParser(description="Attack type x intensity sweep with measured detection outcomes")
    p.add_argument("--model", default="resnet32")
    p.add_argument("--dataset", default="cifar10")
    p.add_argument("--nodes", type=int, default=2)
    p.add_argument("--epochs", type=int, default=3)
    p.add_argument("--steps-per-epoch", type=int, default=10)
    p.add_argument("--attack-start-epoch", type=int, default=2)
    p.add_argument("--batch-size", type=int, default=16)
    p.add_argument("--lr", type=float, default=1e-3)
    p.add_argument("--targets", type=int, nargs="+", default=[1])
    p.add_argument("--types", nargs="+", default=["gradient_poisoning", "output_poisoning", "model_poisoning"])
    p.add_argument("--modes", nargs="+", default=["scale", "sign_flip"])
    p.add_argument("--intensities", type=float, nargs="+", default=[0.05, 0.1, 0.25, 0.5, 1.0])
    p.add_argument("--output-dir", default="results/sweep")
    p.add_argument("--device", default=None)
    p.add_argument("--seq-len", type=int, default=None)
    a = p.parse_args(argv)
    over: Dict[str, Any] = {}
    if a.device:
        over["device"] = a.device
    if a.seq_len:
        over["seq_len"] = a.seq_len
    base = ExperimentConfig(experiment_name="sweep", model_name=a.model, dataset_name=a.dataset, num_nodes=a.nodes, num_epochs=a.epochs,
                            batch_size=a.batch_size, learning_rate=a.lr, attack_start_epoch=a.attack_start_epoch,
                            target_nodes=list(a.targets), steps_per_epoch=a.steps_per_epoch, output_dir=a.output_dir,
                            training_overrides=over)
    rows = attack_sweep(base, a.types, a.intensities, a.modes)
    print((Path(a.output_dir) / "sweep_table.md").read_text())
    return 0 if rows else 1


if __name__ == "__main__":
    logging.basicConfig(level=logging.WARNING)
    raise SystemExit(main())
