"""world_size>1 plumbing on CPU/gloo: BASELINE config 1 (ResNet-32, 2-stage model parallel, one
gradient-poisoning node), 1F1B pipeline == single-process numerics, trust-weighted DP."""
import pytest
import torch

from mp_utils import run_distributed
from trustworthy_dl.parallel.pipeline import make_schedule


def _mp_worker(rank, world, parallelism, micro, attack_step, steps, optimizer="sgd"):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = None
    if attack_step is not None:
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                               start_step=attack_step))
        att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world,
                         learning_rate=1e-3, checkpoint_interval=0, device="cpu", parallelism=parallelism,
                         micro_batches=micro, optimizer=optimizer)
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=steps)]
    opt_steps = {}
    for sid, opt in t.optimizers.items():
        st = [v for v in opt.state_dict()["state"].values()]
        opt_steps[sid] = (int(max(float(v["step"]) for v in st)) if st and "step" in st[0] else None,
                          sorted(st[0].keys()) if st else [])
    return {"losses": losses, "opt_steps": opt_steps, "rx": sorted(t._backend_impl._known_rx), "tx": sorted(t._backend_impl._known_tx),
            "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history],
            "reassign": [(r["from_node"], r["to_node"], r["step"], r["stages"]) for r in t.reassignment_history],
            "owner": list(t.stage_map.owner), "trust": t.trust_manager.trust_vector().tolist(),
            "status": [t.trust_manager.get_node_status(n).value for n in range(world)]}


def test_schedules_are_topological():
    for kind in ("gpipe", "1f1b"):
        for S, M in ((2, 1), (4, 8), (3, 2), (8, 8)):
            order = make_schedule(S, M, kind)
            assert len(order) == 2 * S * M and len(set(order)) == 2 * S * M
            pos = {op: i for i, op in enumerate(order)}
            for s in range(S):
                for m in range(M):
                    if s > 0:
                        assert pos[("F", s - 1, m)] < pos[("F", s, m)]
                    if s < S - 1:
                        assert pos[("B", s + 1, m)] < pos[("B", s, m)]
                    assert pos[("F", s, m)] < pos[("B", s, m)]
    # 1F1B bounds in-flight activations of stage 0 by the number of stages
    order = make_schedule(4, 8, "1f1b")
    live = peak = 0
    for k, s, m in order:
        if s == 0:
            live += 1 if k == "F" else -1
            peak = max(peak, live)
    assert peak <= 4


def test_config1_resnet32_two_stage_gloo_with_poisoned_node():
    res = run_distributed(_mp_worker, 2, "model", 1, 12, 18)
    a, b = res
    assert a["losses"] == pytest.approx(b["losses"])                     # replicated view of the loss
    assert a["attacks"] == b["attacks"] == [(1, 12, "gradient_poisoning")]
    assert a["reassign"] == b["reassign"] == [(1, 0, 12, [1])]
    assert a["owner"] == b["owner"] == [0, 0]
    assert a["trust"] == pytest.approx(b["trust"]) and a["trust"][1] < 0.5 < a["trust"][0]
    assert all(torch.isfinite(torch.tensor(a["losses"])))
    assert len(a["losses"]) == 18                                        # kept training after migration


def test_migration_carries_optimizer_state():
    """The migrated stage continues with the Adam moments and step count it had on the source rank."""
    res = run_distributed(_mp_worker, 2, "model", 1, 12, 18, "adamw")
    a, b = res
    assert a["reassign"] == b["reassign"] == [(1, 0, 12, [1])]
    assert b["opt_steps"] == {}                                          # the source rank gave the optimizer up
    s0, s1 = a["opt_steps"][0], a["opt_steps"][1]
    assert s0[1] == s1[1] and "exp_avg" in s1[1] and "exp_avg_sq" in s1[1]
    assert s0[0] == 18                                                   # stage 0 never moved
    assert s1[0] in (17, 18)                                             # 11 steps on rank 1 + the rest on rank 0 (the
                                                                         # poisoned step's update may be dropped), not 6


def test_model_parallel_matches_single_process():
    from trustworthy_dl import DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=2, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", optimizer="sgd")
    t = DistributedTrainer(cfg)
    single = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=5)]
    multi = run_distributed(_mp_worker, 2, "model", 1, None, 5)[0]["losses"]
    assert multi == pytest.approx(single, rel=1e-4)


def test_pipeline_1f1b_runs_and_learns():
    res = run_distributed(_mp_worker, 2, "pipeline", 4, None, 12)
    assert res[0]["losses"] == pytest.approx(res[1]["losses"])
    # message layouts are learnt from the first header and reused (payload-only messages, no host round trip) afterwards
    assert res[0]["tx"] == res[1]["rx"] == [("F", 0, 2)] and res[1]["tx"] == res[0]["rx"] == [("B", 0, 2)]
    assert res[0]["attacks"] == [] and res[0]["owner"] == [0, 1]


def _dp_worker(rank, world, mode, steps, attack_step):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                           start_step=attack_step, gradient_mode=mode))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=12, num_nodes=world,
                         learning_rate=1e-3, checkpoint_interval=0, device="cpu", parallelism="data", optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 12, num_batches=steps)]
    w = torch.cat([p.detach().reshape(-1) for p in t.model_partitions[0].parameters()])
    return {"losses": losses, "wsum": float(w.double().sum()), "attacks": [(a["node_id"], a["step"]) for a in t.attack_history],
            "owner": list(t.stage_map.owner), "trust": t.trust_manager.trust_vector().tolist(),
            "coeff": t._backend_impl.reducer.last_coeff.tolist()}


@pytest.mark.parametrize("mode", ["scale", "sign_flip"])
def test_data_parallel_trust_weighted_allreduce(mode):
    res = run_distributed(_dp_worker, 3, mode, 10, 6)
    assert len({round(r["wsum"], 6) for r in res}) == 1                  # replicas stay in sync
    assert all(r["attacks"] == res[0]["attacks"] for r in res)
    assert res[0]["attacks"][0] == (1, 6)                                # flagged the step it was injected
    assert res[0]["owner"][1] != 1                                       # its data shard was reassigned
    assert res[0]["coeff"][1] == 0.0                                     # and it no longer contributes
    assert res[0]["trust"][1] < 0.5


def _dp_param_worker(rank, world, steps, attack_step, targets):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(rank)                                              # replicas are synchronised by setup, not by the seed
    att = AdversarialAttacker(AttackConfig(attack_types=["model_poisoning"], target_nodes=targets, intensity=0.5,
                                           start_step=attack_step, end_step=attack_step + 1))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=12, num_nodes=world,
                         learning_rate=1e-3, checkpoint_interval=0, device="cpu", parallelism="data", optimizer="sgd",
                         parameter_monitoring_enabled=True)
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 12, num_batches=steps)]
    w = torch.cat([p.detach().reshape(-1) for p in t.model_partitions[0].parameters()])
    val = t._backend_impl.validate(get_dataloader("cifar10", "validation", 12, num_batches=2))
    return {"losses": losses, "wsum": float(w.double().sum()), "val": val,
            "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history],
            "owner": list(t.stage_map.owner), "trust": t.trust_manager.trust_vector().tolist()}


def test_data_parallel_replica_check_flags_and_repairs_a_poisoned_replica():
    clean = run_distributed(_dp_param_worker, 3, 8, 10 ** 6, [1])
    res = run_distributed(_dp_param_worker, 3, 8, 5, [1])
    assert len({r["wsum"] for r in clean}) == 1 and clean[0]["attacks"] == []   # bit-identical replicas, no false alarm
    assert len({r["val"] for r in clean}) == 1                                  # and one validation loss (BN statistics averaged)
    assert all(r["attacks"] == [(1, 5, "model_poisoning")] for r in res)        # flagged before the poisoned forward ran
    assert len({r["wsum"] for r in res}) == 1                                   # the replica was repaired from the majority
    assert res[0]["losses"][:4] == clean[0]["losses"][:4]
    assert res[0]["owner"][1] != 1 and res[0]["trust"][1] < 0.5


def test_data_parallel_replica_check_without_majority_blames_nobody():
    res = run_distributed(_dp_param_worker, 2, 6, 4, [1])
    assert all(r["attacks"] == [] for r in res)
    assert len({r["wsum"] for r in res}) == 1                                   # still resynchronised


def _tp_worker(rank, world, steps, attack_step):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = None
    if attack_step is not None:
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                               start_step=attack_step))
        att.activate_attacks()
    cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=4, num_nodes=world, learning_rate=1e-2,
                         checkpoint_interval=0, device="cpu", parallelism="tensor", optimizer="sgd", seq_len=32,
                         model_kwargs={"n_positions": 32})
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("tiny-tokens", "train", 4, num_batches=steps, seq_len=32)]
    return {"losses": losses, "attacks": [(a["node_id"], a["step"]) for a in t.attack_history],
            "trust": t.trust_manager.trust_vector().tolist()}


def test_tensor_parallel_matches_single_process():
    from trustworthy_dl import DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=4, num_nodes=1, learning_rate=1e-2,
                         checkpoint_interval=0, device="cpu", optimizer="sgd", seq_len=32, model_kwargs={"n_positions": 32},
                         attack_detection_enabled=False)
    t = DistributedTrainer(cfg)
    single = [t.train_step(b) for b in get_dataloader("tiny-tokens", "train", 4, num_batches=4, seq_len=32)]
    res = run_distributed(_tp_worker, 2, 4, None)
    assert res[0]["losses"] == pytest.approx(res[1]["losses"])
    assert res[0]["losses"] == pytest.approx(single, rel=2e-3)


def test_tensor_parallel_flags_poisoned_rank():
    res = run_distributed(_tp_worker, 2, 6, 3)
    assert res[0]["attacks"] == res[1]["attacks"] and res[0]["attacks"][0] == (1, 3)
    assert res[0]["trust"][1] < 0.5 < res[0]["trust"][0]


def _hybrid_worker(rank, world, attack_step, steps):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = None
    if attack_step is not None:
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[3], intensity=0.5,
                                               start_step=attack_step))
        att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=16, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=2, pp_size=2, dp_size=2,
                         optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 16, num_batches=steps)]
    own = t._backend_impl._my_stages()
    wsum = float(sum(p.detach().double().sum() for s in own for p in t.model_partitions[s].parameters()))
    return {"losses": losses, "attacks": [(a["node_id"], a["step"]) for a in t.attack_history], "owner": list(t.stage_map.owner),
            "reassign": [(r["from_node"], r["to_node"], r["step"]) for r in t.reassignment_history], "own": own, "wsum": wsum}


def test_pipeline_x_data_parallel_hybrid_with_compromise():
    """BASELINE config 5 plumbing (PP x DP): 2 pipeline replicas of 2 stages, rank 3 (replica 1, stage 1) poisons its
    gradients mid-run -> excluded from the DP aggregation in the same step, compromised, its stage migrates to rank 2."""
    res = run_distributed(_hybrid_worker, 4, 10, 14, timeout=400.0)
    assert all(r["losses"] == pytest.approx(res[0]["losses"]) for r in res)
    assert res[0]["attacks"][0] == (3, 10) and res[0]["reassign"][0] == (3, 2, 10)
    assert res[0]["owner"] == [0, 1, 2, 2]
    assert all(torch.isfinite(torch.tensor(res[0]["losses"])))
    # DP replicas of the same stage stayed identical until the compromise (stage 0: ranks 0 and 2 own it)
    assert res[0]["own"] == [0] and 0 in res[2]["own"]


def test_hybrid_replicas_stay_in_sync_without_attack():
    res = run_distributed(_hybrid_worker, 4, None, 6, timeout=400.0)
    assert res[0]["own"] == [0] and res[2]["own"] == [0] and res[0]["wsum"] == pytest.approx(res[2]["wsum"], rel=1e-6)
    assert res[1]["wsum"] == pytest.approx(res[3]["wsum"], rel=1e-6)


# ---------------------------------------------------------------------------------------------------------------------
# bucketed aggregation (what the fused engine runs per gradient segment) against the one-bucket rule
# ---------------------------------------------------------------------------------------------------------------------
def _ranks(W, N, seed=0, noise=0.3):
    import torch
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(N, generator=g)
    return [base + noise * torch.randn(N, generator=g) for _ in range(W)]


@pytest.mark.parametrize("W", [2, 4, 8])
@pytest.mark.parametrize("attack", ["none", "scale", "sign_flip", "nan", "one_bucket_only"])
def test_bucketed_aggregation_verdicts_match_the_single_bucket_rule(W, attack):
    """Per-bucket norms / coefficients (the segments of the overlapped engine) flag the same ranks as one coefficient
    set over the whole gradient for every injected-poisoning scenario, and honest steps give the same aggregate."""
    import torch
    from trustworthy_dl.ops import reference_impl as R
    N = 4096
    bounds = [(0, 1024), (1024, 2048), (2048, 3072), (3072, 4096)]
    gs = _ranks(W, N, seed=W)
    bad = W - 1
    if attack == "scale":
        gs[bad] = gs[bad] * 11.0
    elif attack == "sign_flip":
        gs[bad] = -3.0 * gs[bad]
    elif attack == "nan":
        gs[bad] = gs[bad].clone(); gs[bad][::5] = float("nan")
    elif attack == "one_bucket_only":
        gs[bad] = gs[bad].clone(); gs[bad][1024:2048] *= 40.0          # poison confined to one segment
    tw = torch.full((W,), 1.0 / W)
    out_b, info_b = R.bucketed_trust_reduce(gs, bounds, tw)
    out_1, info_1 = R.bucketed_trust_reduce(gs, [(0, N)], tw)
    assert torch.isfinite(out_b).all() and torch.isfinite(out_1).all()
    if attack == "none":
        assert info_b["flags"].tolist() == info_1["flags"].tolist() == [0] * W
        assert torch.allclose(out_b, out_1, atol=1e-5)
        assert torch.allclose(out_b, torch.stack(gs).mean(0), atol=1e-5)
    elif attack == "sign_flip" and W < 3:
        pass                               # two ranks cannot tell which one flipped (cosine needs a majority); norms decide
    else:
        assert info_b["flags"][bad] == 1 and info_b["flags"][:bad].sum() == 0, info_b
        if attack != "one_bucket_only":
            assert info_1["flags"].tolist() == info_b["flags"].tolist()
    if attack == "one_bucket_only":
        # the per-bucket rule rejects the rank in the poisoned segment only, and still flags it for the step
        assert info_b["rejected"][bad]
        honest = torch.stack(gs[:bad]).mean(0)
        assert torch.allclose(out_b[1024:2048], honest[1024:2048], atol=1e-4)


def _resume_worker(rank, world, ckpt_dir, phase, attack_step):
    """phase 'a': train 14 steps (attack + reassignment at ``attack_step``), checkpoint at step 14, train 4 more.
    phase 'b': fresh trainers resume from the step-14 files and train the same 4 steps."""
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                           start_step=attack_step))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=2, optimizer="adamw",
                         checkpoint_dir=ckpt_dir)
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    batches = list(get_dataloader("cifar10", "train", 8, num_batches=18))
    if phase == "a":
        for b in batches[:14]:
            t.train_step(b)
        t.save_checkpoint()
    else:
        t.load_checkpoint()
    tail = [t.train_step(b) for b in batches[14:]]
    return {"tail": tail, "owner": list(t.stage_map.owner), "opt": sorted(t.optimizers), "step": t.global_step}


def test_multi_process_resume_after_reassignment(tmp_path):
    """Per-rank checkpoint files written after a stage changed hands: resumed ranks take over the saved ownership (modules,
    optimizers with their state) and continue exactly like the run that was not interrupted."""
    a = run_distributed(_resume_worker, 2, str(tmp_path), "a", 10)
    b = run_distributed(_resume_worker, 2, str(tmp_path), "b", 10)
    assert a[0]["owner"] == b[0]["owner"] == [0, 0] and a[0]["opt"] == b[0]["opt"] == [0, 1] and b[1]["opt"] == []
    assert b[0]["step"] == 18
    assert b[0]["tail"] == pytest.approx(a[0]["tail"], rel=1e-5)
    assert b[1]["tail"] == pytest.approx(b[0]["tail"])


def _resume_any_worker(rank, world, ckpt_dir, phase, parallelism):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5, start_step=7))
    att.activate_attacks()
    if parallelism == "tensor":
        cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=4, num_nodes=world, learning_rate=1e-2,
                             checkpoint_interval=0, device="cpu", parallelism="tensor", optimizer="adamw", seq_len=32,
                             model_kwargs={"n_positions": 32}, checkpoint_dir=ckpt_dir)
        batches = list(get_dataloader("tiny-tokens", "train", 4, num_batches=14, seq_len=32))
    else:
        cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=12, num_nodes=world, learning_rate=1e-3,
                             checkpoint_interval=0, device="cpu", parallelism="data", optimizer="adamw", checkpoint_dir=ckpt_dir,
                             parameter_monitoring_enabled=True)
        batches = list(get_dataloader("cifar10", "train", 12, num_batches=14))
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    if phase == "a":
        for b in batches[:10]:
            t.train_step(b)
        t.save_checkpoint()
    else:
        t.load_checkpoint()
    tail = [t.train_step(b) for b in batches[10:]]
    return {"tail": tail, "trust": t.trust_manager.trust_vector().tolist(), "attacks": [(a["node_id"], a["step"]) for a in t.attack_history],
            "status": [t.trust_manager.get_node_status(n).value for n in range(world)]}


@pytest.mark.parametrize("parallelism", ["data", "tensor"])
def test_multi_process_resume_identical_continuation(tmp_path, parallelism):
    """Checkpoint after the attack was handled (rank 1 flagged at step 7), resume in fresh processes: same losses, same trust
    vector, same statuses, and the attack is not re-reported."""
    world = 3 if parallelism == "data" else 2
    a = run_distributed(_resume_any_worker, world, str(tmp_path), "a", parallelism)
    b = run_distributed(_resume_any_worker, world, str(tmp_path), "b", parallelism)
    assert a[0]["attacks"] == b[0]["attacks"] and a[0]["attacks"][0] == (1, 7)
    assert b[0]["tail"] == pytest.approx(a[0]["tail"], rel=1e-5)
    assert b[0]["trust"] == pytest.approx(a[0]["trust"]) and a[0]["status"] == b[0]["status"]


def _train_loop_worker(rank, world, ckpt_dir, parallelism):
    import os
    from trustworthy_dl import DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world, learning_rate=1e-3,
                         num_epochs=2, checkpoint_interval=3, device="cpu", parallelism=parallelism, micro_batches=2,
                         optimizer="sgd", checkpoint_dir=ckpt_dir, lr_schedule="cosine")
    t = DistributedTrainer(cfg)
    t._rank, t._world = rank, world
    t.train(get_dataloader("cifar10", "train", 8, num_batches=4), get_dataloader("cifar10", "validation", 8, num_batches=2))
    stats = t.get_training_stats()
    return {"state": stats["training_state"], "step": stats["global_step"], "files": sorted(os.listdir(ckpt_dir)),
            "lr": [g["lr"] for o in t.optimizers.values() for g in o.param_groups], "loss": stats["metrics"]["final_loss"]}


@pytest.mark.parametrize("parallelism", ["pipeline"])
def test_full_train_loop_multi_process(tmp_path, parallelism):
    """``DistributedTrainer.train`` (epochs, validation, per-rank checkpoints every ``checkpoint_interval`` steps, scheduler
    steps) under two processes."""
    res = run_distributed(_train_loop_worker, 2, str(tmp_path), parallelism)
    for r in res:
        assert r["state"] == "completed" and r["step"] == 8
        assert {"checkpoint_step_3.rank0.pt", "checkpoint_step_3.rank1.pt", "checkpoint_step_6.rank0.pt",
                "checkpoint_step_6.rank1.pt"} <= set(r["files"])
        assert all(lr < 1e-3 for lr in r["lr"])                          # the cosine schedule stepped once per epoch
    assert res[0]["loss"] == pytest.approx(res[1]["loss"])


def _two_bad_worker(rank, world):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1, 2], intensity=0.5, start_step=8))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=2, optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=12)]
    return {"losses": losses, "owner": list(t.stage_map.owner), "attacks": sorted((a["node_id"], a["step"]) for a in t.attack_history),
            "moves": [(r["from_node"], r["to_node"], r["stages"]) for r in t.reassignment_history], "opt": sorted(t.optimizers)}


def test_two_stages_compromised_in_the_same_step():
    """Ranks 1 and 2 of a three-stage pipeline poison their gradients in the same step: both are flagged, both stages move to
    the remaining trusted rank (one after the other, every rank taking the same decisions), training continues there."""
    res = run_distributed(_two_bad_worker, 3)
    for r in res:
        assert r["attacks"] == [(1, 8), (2, 8)] and r["owner"] == [0, 0, 0]
        assert [m[:2] for m in r["moves"]] == [(1, 0), (2, 0)]
        assert r["losses"] == pytest.approx(res[0]["losses"]) and all(torch.isfinite(torch.tensor(r["losses"])))
    assert res[0]["opt"] == [0, 1, 2] and res[1]["opt"] == res[2]["opt"] == []


def _output_poison_worker(rank, world, target):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["output_poisoning"], target_nodes=[target], intensity=0.5, start_step=14))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=2, optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=17)]
    s = t.metrics_collector.get_summary()
    return {"owner": list(t.stage_map.owner), "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history],
            "losses": losses, "det": s.get("detection_rate"), "fp": s.get("false_positive_rate"),
            "status": [t.trust_manager.get_node_status(n).value for n in range(world)]}


@pytest.mark.parametrize("target", [2])
def test_poisoned_stage_output_taints_the_step_not_the_other_nodes(target):
    """A stage emits a corrupted activation.  Everything computed from it in that step -- later stage outputs, the loss
    (it jumps), every gradient (they explode) -- is tainted: only the stage that produced it is flagged, the honest stages
    around it are neither blamed for the exploding gradients nor polluted in their baselines, the step's update is dropped,
    and every process scores the same detection rate from the replicated ground truth."""
    res = run_distributed(_output_poison_worker, 3, target)
    for r in res:
        assert r["attacks"] == [(target, 14, "output_anomaly")]
        assert target not in r["owner"] and r["status"][target] == "compromised"
        assert [st for n, st in enumerate(r["status"]) if n != target] == ["trusted", "trusted"]
        assert r["det"] == 1.0 and r["fp"] == 0.0
        assert r["losses"] == pytest.approx(res[0]["losses"])


def _probation_worker(rank, world, end_step):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5, start_step=8,
                                           end_step=end_step))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=12, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="data", optimizer="sgd")
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    coeff_on_probation = []
    for b in get_dataloader("cifar10", "train", 12, num_batches=34):
        t.train_step(b)
        if not t.stage_map.stages_of(1):
            coeff_on_probation.append(float(t._backend_impl.reducer.last_coeff[1]))
    w = torch.cat([p.detach().reshape(-1) for p in t.model_partitions[0].parameters()])
    return {"owner": list(t.stage_map.owner), "trust": t.trust_manager.get_trust_score(1), "status": t.trust_manager.get_node_status(1).value,
            "moves": [(r["from_node"], r["to_node"], r.get("kind")) for r in t.reassignment_history], "wsum": float(w.double().sum()),
            "probation_coeff": coeff_on_probation, "attacks": [(a["node_id"], a["step"]) for a in t.attack_history]}


def test_probation_a_demoted_rank_earns_its_shard_back_only_when_it_behaves():
    """README 'recovery patterns for nodes after attack mitigation': the demoted rank keeps computing its home shard with
    coefficient 0.  Attack over after three steps -> its measured behaviour lifts its trust back over the threshold and the
    shard returns; attack still running -> it stays excluded (and is re-flagged whenever its status drifts up)."""
    healed = run_distributed(_probation_worker, 3, 11, timeout=400)
    for r in healed:
        assert r["attacks"] == [(1, 8)] and r["owner"] == [0, 1, 2] and r["status"] == "trusted" and r["trust"] > 0.7
        assert [m[2] for m in r["moves"]] == ["data_shard", "shard_returned"] and r["moves"][1][1] == 1
        assert r["probation_coeff"] and max(r["probation_coeff"]) == 0.0       # never influenced the model while on probation
    assert len({r["wsum"] for r in healed}) == 1
    stubborn = run_distributed(_probation_worker, 3, None, timeout=400)
    for r in stubborn:
        assert 1 not in r["owner"] and r["trust"] < 0.5 and max(r["probation_coeff"]) == 0.0
        assert len(r["attacks"]) >= 2 and all(n == 1 for n, _ in r["attacks"])
    assert len({r["wsum"] for r in stubborn}) == 1


def _pp_param_worker(rank, world, attack):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.utils.data_loader import get_dataloader
    torch.manual_seed(0)
    att = None
    if attack:
        att = AdversarialAttacker(AttackConfig(attack_types=["model_poisoning"], target_nodes=[1], intensity=0.5, start_step=9,
                                               end_step=10))
        att.activate_attacks()
    cfg = TrainingConfig(model_name="resnet32", dataset_name="cifar10", batch_size=8, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=2, optimizer="sgd",
                         parameter_monitoring_enabled=True)
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    losses = [t.train_step(b) for b in get_dataloader("cifar10", "train", 8, num_batches=13)]
    return {"losses": losses, "owner": list(t.stage_map.owner),
            "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history]}


def test_pipeline_parameter_monitor_neutralises_model_poisoning():
    """Rank 1's stage is overwritten with noise before step 9.  Its parameter monitor sees parameters that changed outside
    the optimizer, rolls the stage back bit-exactly BEFORE the forward, every rank learns the verdict with the step's
    decision vector and the stage changes hands.  The attacked step itself therefore computes exactly what the clean run
    computes; afterwards the curves differ only by the flagged rank's dropped update of that step."""
    clean = run_distributed(_pp_param_worker, 2, False)
    hit = run_distributed(_pp_param_worker, 2, True)
    assert clean[0]["attacks"] == [] and clean[0]["owner"] == [0, 1]
    for r in hit:
        assert r["attacks"] == [(1, 9, "model_poisoning")] and r["owner"] == [0, 0]
        assert r["losses"][:9] == pytest.approx(clean[0]["losses"][:9], rel=1e-5)        # step 9 ran on the restored weights
        assert r["losses"][9:] == pytest.approx(clean[0]["losses"][9:], rel=3e-2)
