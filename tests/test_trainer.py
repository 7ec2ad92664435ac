"""In-process trainer: API parity, attack -> detection -> trust -> reassignment, checkpoints."""
import os

import pytest
import torch

from trustworthy_dl import (AdversarialAttacker, AttackConfig, DistributedTrainer, NodeStatus, TrainingConfig,
                            TrainingState, TrustManager)
from trustworthy_dl.utils.data_loader import get_dataloader


def _cfg(tmp_path, **kw):
    base = dict(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_epochs=1, num_nodes=2,
                learning_rate=1e-3, checkpoint_interval=0, device="cpu", checkpoint_dir=str(tmp_path / "ck"))
    base.update(kw)
    return TrainingConfig(**base)


def test_readme_constructor_forms():
    t = DistributedTrainer(model_name="gpt2-tiny", num_nodes=3, trust_threshold=0.6, device="cpu")
    assert t.config.num_nodes == 3 and t.trust_manager.trust_threshold == 0.6
    assert t.training_state is TrainingState.INITIALIZING
    with pytest.raises(TypeError):
        DistributedTrainer(TrainingConfig(), model_name="x")


def test_setup_populates_tables(tmp_path):
    t = DistributedTrainer(_cfg(tmp_path, num_nodes=4)).setup()
    assert len(t.model_partitions) == 4 and len(t.optimizers) == 4 and len(t.node_configs) == 4   # Q5
    assert t.node_configs[2].model_partition == "partition_2"


def test_forward_returns_tuple_and_loss_decreases(tmp_path):
    torch.manual_seed(0)
    t = DistributedTrainer(_cfg(tmp_path, learning_rate=3e-3))
    dl = get_dataloader("cifar10", "train", 8, num_batches=30)
    batch = next(iter(dl))
    out, node_outputs = t.forward_pass(batch["input"], [0, 1])            # Q2
    assert out.shape == (8, 10) and set(node_outputs) == {0, 1}
    losses = [t.train_step(b) for b in dl]
    assert sum(losses[-5:]) < sum(losses[:5])
    assert all(v > 0.85 for v in t.get_training_stats()["trust_scores"].values())
    assert t.attack_history == []


def test_gradient_poisoning_detect_compromise_reassign(tmp_path):
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                           start_step=12))
    att.activate_attacks()
    t = DistributedTrainer(_cfg(tmp_path), attacker=att)
    dl = get_dataloader("cifar10", "train", 8, num_batches=20)
    before = None
    for i, b in enumerate(dl):
        if i == 11:
            before = [p.detach().clone() for p in t.model_partitions[1].parameters()]
        t.train_step(b)
        if i == 11:
            after = list(t.model_partitions[1].parameters())
            assert all(torch.equal(a, b_) for a, b_ in zip(before, after))   # poisoned update was dropped
    assert t.attack_history[0]["step"] == 12 and t.attack_history[0]["attack_type"] == "gradient_poisoning"
    assert t.trust_manager.get_node_status(1) in (NodeStatus.COMPROMISED, NodeStatus.SUSPICIOUS)
    r = t.reassignment_history[0]
    assert (r["from_node"], r["to_node"], r["step"]) == (1, 0, 12)
    numel = sum(p.numel() for p in t.model_partitions[1].parameters())
    assert r["migration_time"] == pytest.approx(numel * 4 / 1024 ** 3 + 2.0)  # reference model (:354-365)
    assert t.stage_map.owner == [0, 0]
    assert t.training_state is TrainingState.UNDER_ATTACK
    assert t.trust_manager.get_trust_score(1) < 0.5 < t.trust_manager.get_trust_score(0)
    assert len(t.attack_history) == 1                                      # node 1 owns nothing afterwards


def test_output_poisoning_detected_by_statistics(tmp_path):
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["output_poisoning"], target_nodes=[0], intensity=0.5,
                                           start_step=15))
    att.activate_attacks()
    t = DistributedTrainer(_cfg(tmp_path, gradient_verification_enabled=False), attacker=att)
    for b in get_dataloader("cifar10", "train", 8, num_batches=16):
        t.train_step(b)
    assert [a["attack_type"] for a in t.attack_history] == ["output_anomaly"]
    assert t.attack_history[0]["node_id"] == 0 and t.attack_history[0]["step"] == 15


def test_max_reassignment_attempts(tmp_path):
    t = DistributedTrainer(_cfg(tmp_path, num_nodes=3, max_reassignment_attempts=1)).setup()
    assert t.reassign_node_tasks(1) in (0, 2)
    t.stage_map.owner[1] = 1
    assert t.reassign_node_tasks(1) is None


def test_checkpoint_roundtrip_identical_continuation(tmp_path):
    torch.manual_seed(0)
    dl = list(get_dataloader("cifar10", "train", 8, num_batches=14))
    t = DistributedTrainer(_cfg(tmp_path))
    for b in dl[:10]:
        t.train_step(b)
    path = t.save_checkpoint()
    assert os.path.basename(path) == "checkpoint_step_10.pt"               # reference naming (:461)
    ck = torch.load(path, weights_only=False)
    assert {"epoch", "global_step", "model_partitions", "optimizers", "trust_scores", "attack_history",
            "reassignment_history"} <= set(ck)
    cont = [t.train_step(b) for b in dl[10:]]
    t2 = DistributedTrainer(_cfg(tmp_path))
    t2.load_checkpoint()
    assert t2.global_step == 10
    cont2 = [t2.train_step(b) for b in dl[10:]]
    assert cont == pytest.approx(cont2, rel=1e-5)
    assert t2.trust_manager.trust_vector() == pytest.approx(t.trust_manager.trust_vector())


def test_train_readme_form_and_validation(tmp_path):
    tm = TrustManager(initial_trust=0.9, decay_rate=0.1, recovery_rate=0.05)
    t = DistributedTrainer(model_name="gpt2-tiny", num_nodes=2, trust_threshold=0.7, device="cpu", batch_size=2,
                           seq_len=16, steps_per_epoch=3, checkpoint_interval=0, model_kwargs={"n_positions": 16})
    out = t.train(dataset="tiny-tokens", epochs=1, trust_manager=tm)
    assert t.trust_manager is tm and out["training_state"] == "completed" and out["global_step"] == 3
    assert "val_loss" in out["history"][0] and out["metrics"]["num_batches"] == 3


def test_two_nodes_poison_in_the_same_step_single_process():
    """The reference runner's own scenario (targets [1, 3], experiment_runner.py:93): both are flagged in the step they
    inject, each stage goes to a node that was NOT flagged (load-aware: different targets), no false positive."""
    from trustworthy_dl import AdversarialAttacker, AttackConfig
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1, 3], intensity=0.5, start_step=8))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=4, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=12)]
    assert sorted((a["node_id"], a["step"]) for a in t.attack_history) == [(1, 8), (3, 8)]
    assert set(t.stage_map.owner) == {0, 2} and all(r["to_node"] in (0, 2) for r in t.reassignment_history)
    s = t.metrics_collector.get_summary()
    assert s["detection_rate"] == 1.0 and s["false_positive_rate"] == 0.0
    assert all(l == l for l in losses)


def test_poisoned_last_stage_output_does_not_blame_upstream_nodes():
    """Single process, same rule: the last stage corrupts its output, the loss and all gradients of that step explode, and
    only the last stage's node is flagged (before: every upstream node failed the gradient-norm check in that step)."""
    from trustworthy_dl import AdversarialAttacker, AttackConfig
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["output_poisoning"], target_nodes=[2], intensity=0.5, start_step=14))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=3, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    w_before = None
    for i, b in enumerate(get_dataloader("cifar10", "train", 8, num_batches=16)):
        if i == 13:
            w_before = [p.detach().clone() for p in t.model_partitions[0].parameters()]
        t.train_step(b)
        if i == 13:                                                   # the tainted step's update was dropped everywhere
            assert all(torch.equal(a, b2) for a, b2 in zip(w_before, t.model_partitions[0].parameters()))
    assert [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history] == [(2, 14, "output_anomaly")]
    assert 2 not in t.stage_map.owner


def test_train_saves_a_final_checkpoint_and_resume_continues_with_the_next_epoch(tmp_path):
    kw = dict(model_name="resnet32", dataset_name="cifar10", batch_size=4, num_nodes=2, learning_rate=1e-3, device="cpu",
              optimizer="sgd", checkpoint_interval=100, checkpoint_dir=str(tmp_path), steps_per_epoch=5)
    t = DistributedTrainer(TrainingConfig(num_epochs=1, **kw))
    t.train()
    assert sorted(os.listdir(tmp_path)) == ["checkpoint_step_5.pt"]              # 5 steps < interval: only the final save
    t2 = DistributedTrainer(TrainingConfig(num_epochs=3, **kw))
    t2.load_checkpoint()
    stats = t2.train()
    assert [h["epoch"] for h in stats["history"]] == [1, 2] and stats["global_step"] == 15
    assert "checkpoint_step_15.pt" in os.listdir(tmp_path)


def test_model_poisoning_is_rolled_back_to_the_last_verified_parameters():
    """Single process, parameter monitor on: node 1's stage is overwritten with noise before step 10.  The drift check
    flags it before the forward, the stage is restored from the monitor's last verified snapshot, then it changes hands;
    nobody else is blamed and the loss curve shows no excursion."""
    from trustworthy_dl import AdversarialAttacker, AttackConfig
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["model_poisoning"], target_nodes=[1], intensity=0.5, start_step=10,
                                           end_step=11))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=2, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", optimizer="sgd", parameter_monitoring_enabled=True)
    t = DistributedTrainer(cfg, attacker=att)
    snap, losses = None, []
    for i, b in enumerate(get_dataloader("cifar10", "train", 8, num_batches=13)):
        if i == 9:                                                    # state after step 9 = what step 10 must start from
            snap = [p.detach().clone() for p in t.model_partitions[1].parameters()]
            orig_fwd = t.forward_pass

            def checked_forward(x, *a, **k):                          # the forward of step 10 sees the restored weights
                assert all(torch.equal(p, q) for p, q in zip(snap, t.model_partitions[1].parameters()))
                return orig_fwd(x, *a, **k)
            t.forward_pass = checked_forward
        losses.append(t.train_step(b))
        if i == 9:
            t.forward_pass = orig_fwd
    assert [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history] == [(1, 10, "model_poisoning")]
    assert t.attack_history[0]["parameter_stats"]["restored_to_last_verified"] and t.stage_map.owner == [0, 0]
    assert max(losses[9:]) < max(losses[:9]) + 0.3
