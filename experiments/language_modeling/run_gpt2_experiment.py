#!/usr/bin/env python
"""README.md:105 — language modelling with model poisoning (GPT-2 on openwebtext-shaped synthetic
tokens).  Defaults to a small GPT-2 so it also runs on CPU; pass --config configs/gpt2_distributed.yaml
for the README's 8-node GPT-2-medium setup."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from trustworthy_dl.experiments.runner import ExperimentConfig, ExperimentRunner, main

if __name__ == "__main__":
    if "--config" in sys.argv:
        raise SystemExit(main(sys.argv[1:]))
    cfg = ExperimentConfig(
        experiment_name="gpt2_language_modeling", model_name="gpt2-tiny", dataset_name="tiny-tokens", num_nodes=4,
        num_epochs=4, batch_size=8, learning_rate=1e-3, attack_enabled=True, attack_start_epoch=2,
        attack_types=["gradient_poisoning", "model_poisoning"], target_nodes=[1, 3], steps_per_epoch=10,
        training_overrides={"seq_len": 64, "parameter_monitoring_enabled": True})
    runner = ExperimentRunner(cfg)
    res = runner.run_experiment()
    print("summary:", res["experiment_summary"])
