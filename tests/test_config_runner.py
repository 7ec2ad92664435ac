import json
import os

import pytest
import yaml

from trustworthy_dl.config import dump_training_config, load_training_config, training_config_from_dict
from trustworthy_dl.experiments.runner import ExperimentConfig, ExperimentRunner, main as runner_main

README_YAML = """
model:
  name: "gpt2"
  size: "medium"
training:
  batch_size: 32
  learning_rate: 5e-5
  num_epochs: 10
distributed:
  num_nodes: 8
  parallelism: "model"
security:
  trust_threshold: 0.7
  attack_detection: true
  gradient_verification: true
"""


def test_readme_yaml_schema(tmp_path):
    p = tmp_path / "gpt2_distributed.yaml"
    p.write_text(README_YAML)
    cfg = load_training_config(str(p))
    assert (cfg.model_name, cfg.model_size, cfg.batch_size, cfg.num_nodes) == ("gpt2", "medium", 32, 8)
    assert cfg.learning_rate == pytest.approx(5e-5) and cfg.parallelism == "model"
    assert cfg.attack_detection_enabled and cfg.gradient_verification_enabled
    again = training_config_from_dict(yaml.safe_load(dump_training_config(cfg)))
    assert again.model_size == "medium" and again.num_nodes == 8
    with pytest.raises(ValueError):
        training_config_from_dict({"distributed": {"parallelism": "zigzag"}})


def test_shipped_configs_parse():
    root = os.path.join(os.path.dirname(os.path.dirname(__file__)), "configs")
    files = [f for f in os.listdir(root) if f.endswith(".yaml")]
    assert len(files) >= 5
    for f in files:
        load_training_config(os.path.join(root, f))


def test_runner_end_to_end_artifacts(tmp_path):
    cfg = ExperimentConfig(experiment_name="t", model_name="resnet32", dataset_name="cifar10", num_nodes=2,
                           num_epochs=3, batch_size=4, learning_rate=1e-3, attack_enabled=True, attack_start_epoch=2,
                           attack_types=["gradient_poisoning"], target_nodes=[1], steps_per_epoch=6,
                           output_dir=str(tmp_path), training_overrides={"device": "cpu"})
    res = ExperimentRunner(cfg).run_experiment()
    out = tmp_path / "t"
    for f in ("experiment_results.json", "training_metrics.csv", "experiment_report.md", "chart_data.json",
              "trust_data.json", "detection_data.json"):
        assert (out / f).exists(), f
    s = res["experiment_summary"]
    assert s["total_batches"] == 18 and s["total_attacks_detected"] >= 1 and s["reassignments"] >= 1
    assert res["final_attack_statistics"]["by_type"]["gradient_poisoning"] >= 1
    data = json.loads((out / "chart_data.json").read_text())
    assert len(data["trust_evolution"]["nodes"]["1"]) == 3
    assert data["trust_evolution"]["nodes"]["1"][-1] < data["trust_evolution"]["nodes"]["0"][-1]  # measured, not simulated
    so = s["security_outcomes"]                               # README "key findings", measured from this run's histories
    assert so["attacked_nodes"] == [1] and so["detected_nodes"] == [1] and so["missed_nodes"] == [] and so["false_positive_nodes"] == []
    assert so["detection_latency_steps"] == {"1": 0} and so["detection_rate"] == 1.0        # flagged the step it was injected
    assert so["reassignment_time_estimated_s"] > 0 and so["loss_before_attack"] > 0
    assert "Measured attack outcomes" in (out / "experiment_report.md").read_text()
    m = res["training_results"]["training_metrics"][-1]
    det = res["detection_statistics"]
    assert det["true_positive_rate"] > 0 and det["false_positive_rate"] == 0     # counters fed from the injector's ground truth


def test_runner_cli_honours_config(tmp_path):
    y = tmp_path / "c.yaml"
    y.write_text("model: {name: resnet32}\ntraining: {batch_size: 4, learning_rate: 0.001, num_epochs: 1, dataset: cifar10}\n"
                 "distributed: {num_nodes: 2, device: cpu}\nexperiment: {name: fromyaml, steps_per_epoch: 2}\n")
    assert runner_main(["--config", str(y), "--output-dir", str(tmp_path)]) == 0          # Q17
    assert (tmp_path / "fromyaml" / "experiment_report.md").exists()


def test_report_with_empty_summary_does_not_raise(tmp_path):
    cfg = ExperimentConfig("e", "resnet32", "cifar10", 2, 1, 4, 1e-3, output_dir=str(tmp_path))
    text = ExperimentRunner(cfg)._create_report_content({"experiment_summary": {}})            # Q18
    assert "N/A" in text


def test_attack_sweep_tabulates_measured_outcomes(tmp_path):
    from trustworthy_dl.experiments.sweep import attack_sweep
    base = ExperimentConfig(experiment_name="s", model_name="resnet32", dataset_name="cifar10", num_nodes=2, num_epochs=3,
                            batch_size=4, learning_rate=1e-3, attack_start_epoch=2, target_nodes=[1], steps_per_epoch=6,
                            output_dir=str(tmp_path), training_overrides={"device": "cpu"})
    rows = attack_sweep(base, ["gradient_poisoning"], [0.5], ["scale"])
    assert [r["attack_type"] for r in rows] == ["none", "gradient_poisoning"]
    assert rows[0]["attacks_detected"] == 0 and rows[0]["false_positive_nodes"] == []          # clean run: nothing flagged
    assert rows[1]["detection_rate"] == 1.0 and rows[1]["mean_detection_latency_steps"] == 0.0 and rows[1]["missed_nodes"] == []
    table = (tmp_path / "sweep_table.md").read_text()
    assert "scale" in table and (tmp_path / "sweep_results.json").exists()
