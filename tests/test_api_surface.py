"""Small public methods that nothing else in the suite calls: reference-parity conveniences (trust_manager.py:294-300,
391-394; attack_detector.py:442-445) and the extensions that feed the README's detection metrics (README.md:158-165)."""
import numpy as np
import torch

from trustworthy_dl import (AdversarialAttacker, AttackConfig, AttackDetector, DistributedTrainer, GradientVerifier,
                            ParameterMonitor, TrainingConfig, TrustManager)
from trustworthy_dl.core.node_monitor import NodeMonitor
from trustworthy_dl.utils.data_loader import get_dataloader


def test_trust_manager_conveniences():
    tm = TrustManager(num_nodes=3, trust_threshold=0.7)
    v = tm.update_all([0.0, 0.5, 1.0], [1.0, 1.0, 0.0])
    assert v.shape == (3,) and v[0] > v[1] > v[2]
    assert v.tolist() == [tm.get_trust_score(i) for i in range(3)]
    for _ in range(3):
        tm.update_trust_score(2, 1.0, 0.0)
    h = tm.get_node_history(2, limit=2)
    assert len(h) == 2 and h[-1]["trust_score"] == tm.get_trust_score(2) and tm.get_node_history(99) == []
    tm.mark_compromised(2, "test")
    tm.reset_node_trust(2)
    assert tm.get_trust_score(2) == 1.0 and tm.get_node_status(2).value == "trusted"


def test_detector_and_verifier_conveniences():
    d = AttackDetector()
    d.set_detection_threshold(1.7)
    assert d.detection_threshold == 1.0
    d.record_ground_truth(True, True); d.record_ground_truth(False, True); d.record_ground_truth(True, False)
    assert d.detection_stats["true_positives"] == 1 and d.detection_stats["false_positives"] == 1
    g = GradientVerifier(clip_norm=2.0)
    assert g.clip_factor(1.0) == 1.0 and g.clip_factor(8.0) == 0.25 and GradientVerifier().clip_factor(1e9) == 1.0
    nm = NodeMonitor()
    nm.record_latency(1, 0.25)
    assert nm.latency[1] == 0.25


def test_parameter_monitor_publish_and_verify():
    pm = ParameterMonitor()
    w = [torch.arange(12, dtype=torch.float32).reshape(3, 4), torch.ones(5)]
    cs = pm.publish(0, w)
    assert pm.verify_received(0, [t.clone() for t in w]) and pm.verify_received(0, w, expected=cs)
    w[1][2] += 1e-3
    assert not pm.verify_received(0, w) and pm.events[-1]["reason"] == "checksum_mismatch"
    assert pm.verify_received(7, w)                                    # nothing published for that shard: nothing to compare


def test_attacker_log_queries_and_detection_rates_from_ground_truth():
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5, start_step=8))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=4, num_nodes=2, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    for b in get_dataloader("cifar10", "train", 4, num_batches=10):
        t.train_step(b)
    assert att.was_attacked(8, 1) and not att.was_attacked(8, 0) and not att.was_attacked(7) and att.was_attacked(8)
    s = t.metrics_collector.get_summary()
    assert s["detection_rate"] == 1.0 and s["false_positive_rate"] == 0.0     # one positive (node 1, step 8), flagged; no false alarm
    assert t.attack_detector.detection_stats["true_positives"] == 1
    att.deactivate_attacks()
    assert not att.is_active() and not att._armed("gradient_poisoning", 9, 1)


def test_model_follows_the_named_dataset():
    """``vgg11`` + ``cifar10`` used to die in the classifier (ImageNet-sized flatten on a 32 x 32 input)."""
    t = DistributedTrainer(TrainingConfig(model_name="vgg11", dataset_name="cifar10", batch_size=4, num_nodes=2, device="cpu",
                                          checkpoint_interval=0))
    loss = t.train_step(next(iter(get_dataloader("cifar10", "train", 4, num_batches=1))))
    assert loss == loss and t.model.classifier[-1].fc.out_features == 10
    t2 = DistributedTrainer(TrainingConfig(model_name="resnet32", dataset_name="imagenet", num_nodes=2, device="cpu"))
    assert t2._dataset_model_kwargs() == {"num_classes": 1000}
    t3 = DistributedTrainer(TrainingConfig(model_name="vgg16", dataset_name="cifar10", num_nodes=2, device="cpu",
                                           model_kwargs={"num_classes": 7}))
    assert t3._dataset_model_kwargs() == {"num_classes": 7, "input_size": 32}            # explicit kwargs win


def test_user_supplied_sequential_model_is_partitioned_by_its_children():
    import torch.nn as nn
    torch.manual_seed(0)
    m = nn.Sequential(nn.Flatten(), nn.Linear(3 * 32 * 32, 64), nn.ReLU(), nn.Linear(64, 64), nn.ReLU(), nn.Linear(64, 10))
    t = DistributedTrainer(TrainingConfig(model_name="custom", dataset_name="cifar10", batch_size=8, num_nodes=3, device="cpu",
                                          checkpoint_interval=0, learning_rate=1e-2), model=m)
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=8)]
    assert t.stage_map.owner == [0, 1, 2] and sum(len(p) for p in t.model_partitions.values()) == 6
    assert losses[-1] < losses[0]


def test_metric_summary_ignores_the_non_finite_loss_of_a_tainted_step():
    from trustworthy_dl.utils.metrics import MetricsCollector
    m = MetricsCollector()
    for i, l in enumerate([2.0, float("nan"), 1.0]):
        m.collect_batch_metrics({"loss": l, "step": i + 1, "epoch": 0, "trust_scores": {0: 1.0}, "step_time": 0.1, "samples": 4})
    s = m.get_summary()
    assert s["mean_loss"] == 1.5 and s["final_loss"] == 1.0 and s["non_finite_losses"] == 1
