"""Host side of the fused data-parallel and tensor-parallel trainer paths (``DataParallelStrategy._engine_step``,
``TensorParallelStrategy._engine_step``: mirroring the device-resident trust state into the public API, attack / reassignment
histories, the ZeRO-1 parameter-monitor verdicts, detection scoring, short batches, checkpoints) executed on CPU ranks over gloo
with a stand-in for the engine.  The stand-in plays the device: it sees every rank's fault-injection inputs (like the real
reduction sees every rank's bucket) and keeps a replicated trust vector / status / flag state with the semantics of K3."""
import pytest
import torch

from mp_utils import run_distributed

T = 128


class _Reducer:
    def __init__(self, world):
        self.trust = torch.ones(world)
        self.status = torch.zeros(world, dtype=torch.int32)


class _Engine:
    """What the strategies touch of GPT2Engine / GPT2TPEngine."""

    def __init__(self, world, rank, zero=0, tokens=0):
        import torch.distributed as dist
        self.dist, self.world, self.rank, self.zero = dist, world, rank, zero
        self.reducer = _Reducer(world)
        self.flags = torch.zeros(world, dtype=torch.int32)
        self.pflags = torch.zeros(world, dtype=torch.int32)
        self.param_attack = torch.zeros(1)
        self.loss_sum = torch.zeros(1)
        self.M = max(1, tokens)
        self.steps = 0
        self.device = torch.device("cpu")

    def train_step(self, idx, tgt, attack_scale=1.0):
        assert idx.shape[-1] == T
        mine = torch.tensor([float(attack_scale), float(self.param_attack[0])], dtype=torch.float64)
        table = torch.zeros(self.world * 2, dtype=torch.float64)
        self.dist.all_gather_into_tensor(table, mine)                    # the reduction sees every rank's bucket
        table = table.view(self.world, 2)
        self.steps += 1
        r = self.reducer
        for n in range(self.world):
            bad = abs(float(table[n, 0]) - 1.0) > 1e-9
            self.flags[n] = 1 if bad else 0
            self.pflags[n] = 1 if float(table[n, 1]) > 0 else 0
            if bad:
                r.trust[n], r.status[n] = 0.1, 2
            elif int(r.status[n]) != 2:
                r.trust[n] = 0.9 * r.trust[n] + 0.1 * 0.9
        loss = 3.0 / (1.0 + 0.1 * self.steps)
        self.loss_sum[0] = loss * self.M / self.world                    # TP: per-rank partial sum
        return loss

    def security_snapshot(self):
        r = self.reducer
        snap = {"trust": r.trust.clone(), "status": r.status.clone(), "flags": self.flags.clone(),
                "stats": torch.ones(self.world, 4), "coeff": torch.ones(self.world), "act_stats": torch.zeros(12, dtype=torch.float64)}
        if self.zero:
            snap["param_flags"] = self.pflags.clone()
            snap["param_drift"] = torch.full((self.world,), 0.01, dtype=torch.float64)
        return snap

    def export_to_module(self, model):
        pass

    def state_dict(self):
        return {"steps": torch.tensor(self.steps), "trust": self.reducer.trust.clone(), "status": self.reducer.status.clone()}

    def load_state_dict(self, sd):
        self.steps = int(sd["steps"])
        self.reducer.trust.copy_(sd["trust"]); self.reducer.status.copy_(sd["status"])


def _trainer(rank, world, parallelism, zero, ckpt_dir):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    from trustworthy_dl.parallel.data_parallel import DataParallelStrategy
    from trustworthy_dl.parallel.tensor_parallel import TensorParallelStrategy
    types = ["model_poisoning"] if zero else ["gradient_poisoning"]      # ZeRO-1 run: the K9 verdict path instead of K2's
    att = AdversarialAttacker(AttackConfig(attack_types=types, target_nodes=[1], intensity=0.5, start_step=5, end_step=6))
    att.activate_attacks()
    cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=2 * world, num_nodes=world, learning_rate=1e-3,
                         checkpoint_interval=0, device="cpu", parallelism=parallelism, optimizer="adamw", seq_len=T, engine="fused",
                         zero=zero, checkpoint_dir=ckpt_dir, model_kwargs={"n_positions": T, "n_embd": 128, "n_head": 2, "n_layer": 2})
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world

    if parallelism == "data":
        class Strategy(DataParallelStrategy):
            def _make_engine(self, model):
                return _Engine(self.world, self.rank, zero=zero)
    else:
        class Strategy(TensorParallelStrategy):
            def _fused_blockers(self, base):
                return []

            def _make_engine(self, base):
                return _Engine(self.world, self.rank, tokens=2 * self.world * T)
    t._backend_impl = Strategy(t)
    t._backend_impl.setup()
    return t, att


def _worker(rank, world, parallelism, zero, ckpt_dir):
    from trustworthy_dl.utils.data_loader import get_dataloader
    t, att = _trainer(rank, world, parallelism, zero, ckpt_dir)
    batches = list(get_dataloader("tiny-tokens", "train", 2 * world, num_batches=8, seq_len=T))
    losses = [t.train_step(b) for b in batches[:7]]
    short = {k: v[:world] for k, v in batches[7].items()}                    # a last batch without drop_last
    losses.append(t.train_step(short))
    eng = t._backend_impl.engine
    out = {"losses": losses, "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history],
           "moves": [(r["from_node"], r.get("kind")) for r in t.reassignment_history], "state": t.training_state.value,
           "status": [t.trust_manager.get_node_status(n).value for n in range(world)],
           "trust": t.trust_manager.trust_vector().tolist(), "dev_trust": eng.reducer.trust.tolist(),
           "tm_attacks": [len(t.trust_manager.attack_history[n]) for n in range(world)], "engine_steps": eng.steps,
           "summary": {k: t.metrics_collector.get_summary().get(k) for k in ("detection_rate", "false_positive_rate", "num_batches")}}
    if parallelism == "data":
        out["val"] = t.validate(get_dataloader("tiny-tokens", "validation", 2 * world, num_batches=1, seq_len=T))
        t.save_checkpoint()
        t2, _ = _trainer(rank, world, parallelism, zero, ckpt_dir)
        t2.load_checkpoint()
        out["resumed"] = {"step": t2.global_step, "engine_steps": t2._backend_impl.engine.steps,
                          "status": [t2.trust_manager.get_node_status(n).value for n in range(world)],
                          "attacks": len(t2.attack_history)}
    return out


@pytest.mark.parametrize("parallelism,zero", [("data", 0), ("data", 1), ("tensor", 0)])
def test_fused_trainer_paths_mirror_the_device_state(tmp_path, parallelism, zero):
    world = 3 if parallelism == "data" else 2
    res = run_distributed(_worker, world, parallelism, zero, str(tmp_path))
    want = [(1, 5, "model_poisoning" if zero else "gradient_poisoning")]
    for r in res:
        assert r["attacks"] == want                                          # recorded once, in the step the device flagged it
        assert r["status"][1] == "compromised" and all(s == "trusted" for n, s in enumerate(r["status"]) if n != 1)
        assert r["trust"] == pytest.approx(r["dev_trust"], abs=1e-6)         # the device-resident vector is authoritative
        assert r["trust"][1] == pytest.approx(0.1) and r["dev_trust"][1] == pytest.approx(0.1) and r["state"] == "under_attack"
        assert r["moves"] and r["moves"][0][0] == 1 and r["tm_attacks"][1] >= 1
        assert r["summary"]["detection_rate"] == 1.0 and r["summary"]["false_positive_rate"] == 0.0
        assert r["engine_steps"] == 7 and r["summary"]["num_batches"] == 7   # the short batch was skipped ...
        assert r["losses"][7] == r["losses"][6]                              # ... and reported the last loss
        assert r["losses"][:7] == pytest.approx(res[0]["losses"][:7])
        if parallelism == "data":
            assert r["val"] == r["val"] and r["resumed"] == {"step": 8, "engine_steps": 7, "status": r["status"], "attacks": len(want)}
