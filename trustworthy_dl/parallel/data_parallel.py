"""Data parallelism with the fused trust-weighted all-reduce (BASELINE config 3 and the headline
GPT-2-medium metric).  The reference imports DDP and never uses it (distributed_trainer.py:9,
SURVEY 2.4).

Every rank is a trust "node" holding a full replica.  Gradients are written straight into one
flat bucket (``param.grad`` are views, no pack copy on the eager path); the bucket is aggregated
by :class:`TrustAllReduce`; the statistics it emits feed the verifier and the trust state machine
identically on every rank, so decisions are replicated without extra traffic.  A rank that is
COMPROMISED gets coefficient 0 and its data shard is handed to the most trusted, least loaded
rank (the DP form of "task reassignment").

On a B200 with a GPT-2 model the replica is executed by the fused engine
(:mod:`trustworthy_dl.engine.gpt2_engine`) instead of eager autograd.
"""
from __future__ import annotations

import logging
import time
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn as nn

from ..utils.timing import nvtx_range
from ..core.partition import StageMap
from . import comm
from .trust_allreduce import TrustAllReduce

logger = logging.getLogger(__name__)

__all__ = ["DataParallelStrategy", "flatten_grads_into"]


def flatten_grads_into(params: List[nn.Parameter], dtype: torch.dtype, device: torch.device,
                       buffer: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Make every ``p.grad`` a view into one flat buffer (allocated here unless given)."""
    total = sum(p.numel() for p in params)
    if buffer is None:
        buffer = torch.zeros(total, dtype=dtype, device=device)
    off = 0
    for p in params:
        n = p.numel()
        p.grad = buffer[off:off + n].view_as(p)
        off += n
    return buffer


class DataParallelStrategy:
    def __init__(self, trainer):
        self.t = trainer
        self.rank, self.world = comm.rank(), comm.world_size()
        self.device = trainer.device
        self.engine = None
        self.flat_params = None
        self._last_loss = float("nan")
        self._sec_s = 0.0
        self._step_s = 0.0

    def setup(self) -> None:
        t, c = self.t, self.t.config
        if c.num_nodes != self.world:
            c.num_nodes = self.world
            t.trust_manager.resize(self.world)
        model = t._build_model().to(self.device)
        t.model_partitions[0] = model
        use_fused = c.engine == "fused" or (c.engine == "auto" and self.device.type == "cuda"
                                            and c.model_name.startswith("gpt2") and c.dtype == "bf16")
        if use_fused:
            # the fused engine runs AdamW at a constant learning rate over full, 128-aligned batches; anything else
            # goes through the eager path (with a log line) instead of being silently ignored
            T = c.seq_len or model.config.n_positions
            why = [m for ok, m in [(T % 128 == 0, f"seq_len {T} is not a multiple of 128"),
                                   (str(c.optimizer).lower() == "adamw", f"optimizer {c.optimizer!r} (engine: adamw)"),
                                   (str(c.lr_schedule).lower() in ("constant", "none", ""), f"lr_schedule {c.lr_schedule!r}"),
                                   (not c.clip_norm, "clip_norm > 0 (the engine clips per rank inside the aggregation)"),
                                   (c.batch_size % self.world == 0, "batch_size not divisible by the world size"),
                                   (model.config.n_embd // model.config.n_head == 64, "head_dim != 64")] if not ok]
            if why:
                if c.engine == "fused":
                    raise ValueError("engine: fused cannot honour this configuration: " + "; ".join(why))
                logger.info("engine: auto -> eager path (%s)", "; ".join(why))
                use_fused = False
        if use_fused:
            self.engine = self._make_engine(model)
            self.reducer = self.engine.reducer
            t.model_partitions[0] = model
        else:
            self.params = [p for p in model.parameters() if p.requires_grad]
            gdtype = torch.float32
            self.reducer = TrustAllReduce(sum(p.numel() for p in self.params), gdtype, self.device)
            flatten_grads_into(self.params, gdtype, self.device, self.reducer.buffer)
            self.flat_params = None
            if c.parameter_monitoring_enabled:
                # one flat parameter vector (every ``p.data`` a view into it, before the optimizer is built) that all
                # replicas start from bit-identically: the per-step replica check reads it with one checksum kernel
                self.flat_params = torch.cat([p.detach().reshape(-1) for p in self.params])
                if self.world > 1:
                    comm.broadcast_tensor(self.flat_params, 0)
                off = 0
                for p in self.params:
                    p.data = self.flat_params[off:off + p.numel()].view_as(p)
                    off += p.numel()
            t.optimizers[0] = t._make_optimizer(self.params)
            t.schedulers[0] = t._make_scheduler(t.optimizers[0])
        # data shards play the role of stages: shard i starts on rank i
        t.stage_map = StageMap(self.world, self.world)
        t.stage_costs = [1.0] * self.world

    def _make_engine(self, model):
        """The fused replica engine (one place, so that tests/test_fused_wrappers_cpu.py can run the host side of
        ``_engine_step`` over gloo with a stand-in)."""
        from ..engine.gpt2_engine import GPT2Engine
        c = self.t.config
        return GPT2Engine.from_module(model, batch_size=max(1, c.batch_size // self.world),
                                      seq_len=c.seq_len or model.config.n_positions,
                                      lr=c.learning_rate, weight_decay=c.weight_decay, device=self.device,
                                      security=bool(c.attack_detection_enabled or c.gradient_verification_enabled),
                                      zero=int(getattr(c, "zero", 0) or 0),
                                      trust_threshold=c.trust_threshold)

    # ------------------------------------------------------------------ step
    def _local_batch(self, batch: Dict[str, torch.Tensor]) -> Optional[Dict[str, torch.Tensor]]:
        shards = self.t.stage_map.stages_of(self.rank)
        if not shards:
            return None
        xs, ys = batch["input"].chunk(self.world), batch["target"].chunk(self.world)
        return {"input": torch.cat([xs[s] for s in shards]), "target": torch.cat([ys[s] for s in shards])}

    def train_step(self, batch: Dict[str, torch.Tensor]) -> float:
        t, c = self.t, self.t.config
        t0 = time.perf_counter()
        step = t.global_step
        if self.engine is not None:
            return self._engine_step(batch, t0)
        model = t.model_partitions[0]
        if self.flat_params is not None:
            self._check_replicas(step)
        local = self._local_batch(batch)
        # Probation ("recovery patterns for nodes after attack mitigation", README.md:141): a rank that lost its data shard
        # keeps computing the shard it started with.  Its gradient enters the reduction with coefficient 0 -- it cannot
        # influence the model -- but its norm, its cosine to the median and its output statistics are measured like
        # everybody's, which is the only evidence a demoted rank can offer; without it its trust saturates at 0.3 for good.
        shadow = local is None and c.probation and t.trust_manager.get_node_status(self.rank).value != "offline"
        if shadow:
            local = {"input": batch["input"].chunk(self.world)[self.rank], "target": batch["target"].chunk(self.world)[self.rank]}
        self.reducer.buffer.zero_()
        loss_val, n_local = 0.0, 0
        out_stats = None
        if local is not None:
            with t._autocast():
                out = model(local["input"].to(self.device))
            if t.attacker is not None:
                out = t.attacker.poison_output(self.rank, out, step)
            loss = t.calculate_loss(out, local["target"])
            flatten_grads_into(self.params, self.reducer.buffer.dtype, self.device, self.reducer.buffer)
            loss.backward()
            loss_val, n_local = float(loss.item()), (0 if shadow else local["input"].shape[0])
            if c.attack_detection_enabled:
                from ..ops.stats import tensor_statistics
                out_stats = tensor_statistics(out.detach()).cpu()
        if t.attacker is not None:
            t.attacker.poison_gradients(self.rank, [self.reducer.buffer], step)
        ts = time.perf_counter()
        w_np = t.trust_manager.aggregation_weights()
        idle = [n for n in range(self.world) if not t.stage_map.stages_of(n)]
        if idle and len(idle) < self.world:                  # a rank on probation computes a shard somebody else covers: weight 0
            w_np = w_np.copy()
            w_np[idle] = 0.0
            w_np = w_np / w_np.sum() if w_np.sum() > 0 else np.full(self.world, 1.0 / self.world)
        weights = torch.from_numpy(w_np).float()
        with nvtx_range("dp.trust_allreduce"):
            stats, coeff = self.reducer.reduce_(weights)
        stats = stats.detach().cpu().double().numpy()
        # ---- replicated security decisions -----------------------------------------
        vec = torch.cat([out_stats if out_stats is not None else torch.zeros(12, dtype=torch.float64),
                         torch.tensor([1.0 if out_stats is not None else 0.0, loss_val, float(n_local)],
                                      dtype=torch.float64)])
        table = comm.all_gather_vector(vec).numpy()
        t._last_step_info = {"out_stats": {}, "invalid_nodes": set(), "grad_norms": {}}
        offenders: List[int] = []                        # judged first, demoted together, data shards reassigned afterwards
        for node in range(self.world):
            if table[node, 12] > 0:
                t._last_step_info["out_stats"][node] = (float(table[node, 0]), float(table[node, 1]))
                if c.attack_detection_enabled:
                    res = t.attack_detector.observe_stats(table[node, :12], node, step)
                    if res.is_attack:
                        t.handle_detected_attack(node, None, result=res, reassign=False)
                        offenders.append(node)
        if c.gradient_verification_enabled:
            ok = t.gradient_verifier.verify_rank_stats(stats[:, 0], stats[:, 1], float(stats[0, 2]), step)
            for node in np.nonzero(~ok)[0]:
                node = int(node)
                t._last_step_info["invalid_nodes"].add(node)         # evidence for the trust score whatever the status
                if t.trust_manager.get_node_status(node).value != "compromised":
                    t.handle_gradient_attack(node, [], reassign=False)
                    offenders.append(node)
        for node in offenders:
            t.reassign_node_tasks(node)
        flagged = {a["node_id"] for a in t.attack_history if a["step"] == step}
        cos = stats[:, 3]
        for node in range(self.world):
            st = t._last_step_info["out_stats"].get(node)
            dev = 1.0 if st is None else self._deviation(node, st)
            gc = self._consistency(cos, node, node in t._last_step_info["invalid_nodes"])
            t.trust_manager.update_trust_score(node, dev, gc)
            if st is not None and node not in flagged:
                t.node_monitor.record_output(node, *st)
        if c.probation:
            self._return_shards(step)
        self._sec_s += time.perf_counter() - ts
        # ---- optimizer on the aggregated gradient ------------------------------------
        opt = t.optimizers[0]
        if c.clip_norm > 0:
            nn.utils.clip_grad_norm_(self.params, c.clip_norm)
        opt.step()
        tot_n = table[:, 14].sum()
        loss = float((table[:, 13] * table[:, 14]).sum() / max(1.0, tot_n))
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss, "step": step, "epoch": t.current_epoch,
            "trust_scores": {i: t.trust_manager.get_trust_score(i) for i in range(self.world)},
            "step_time": time.perf_counter() - t0, "samples": batch["input"].shape[0]})
        return loss

    def _engine_step(self, batch, t0) -> float:
        """Fused path: the whole step (fwd, bwd, statistics, trust-weighted reduction, trust state machine,
        AdamW) runs on the device; afterwards the host trust manager / histories are refreshed from the
        device-resident state (a few hundred bytes) so the public API keeps reporting what happened."""
        from ..core.trust_manager import NodeStatus, STATUS_CODES
        t, c = self.t, self.t.config
        step = t.global_step
        if comm.is_short_batch(batch, c.batch_size):
            return self._last_loss
        # static shapes on this path: every rank always processes its own slice; a compromised rank simply gets
        # coefficient 0 in the reduction (its shard is recorded as reassigned)
        local = {"input": batch["input"].chunk(self.world)[self.rank], "target": batch["target"].chunk(self.world)[self.rank]}
        scale = 1.0
        if t.attacker is not None and t.attacker._armed("gradient_poisoning", step, self.rank):
            i = t.attacker.config.intensity
            scale = -(1.0 + 4.0 * i) if t.attacker.config.gradient_mode == "sign_flip" else 1.0 + 20.0 * i
            t.attacker._record("gradient_poisoning", step, self.rank, mode=t.attacker.config.gradient_mode)
        if self.engine.zero and t.attacker is not None:
            armed = t.attacker._armed("model_poisoning", step, self.rank)
            self.engine.param_attack.fill_(float(t.attacker.config.intensity) / 10.0 if armed else 0.0)
            if armed:
                t.attacker._record("model_poisoning", step, self.rank)
        loss = self.engine.train_step(local["input"], local["target"], attack_scale=scale)
        ts = time.perf_counter()
        snap = self.engine.security_snapshot()
        trust, status, flags = snap["trust"].tolist(), snap["status"].tolist(), snap["flags"].tolist()
        tm = t.trust_manager
        for node in range(self.world):
            if flags[node] & 1 and tm.get_node_status(node) is not NodeStatus.COMPROMISED:
                t.attack_history.append({"node_id": node, "timestamp": time.time(), "step": step,
                                         "attack_type": "gradient_poisoning",
                                         "gradient_stats": {"total_norm": float(snap["stats"][node, 0]),
                                                            "cosine_to_median": float(snap["stats"][node, 3])}})
                tm.attack_history[node].append({"timestamp": tm.clock(), "attack_type": "gradient_poisoning",
                                                "previous_trust": tm.get_trust_score(node)})
                t.metrics_collector.record_event("attack", node_id=node, type="gradient_poisoning", step=step)
                self.reassign(node)
                from ..core.distributed_trainer import TrainingState
                t.training_state = TrainingState.UNDER_ATTACK
            tm._value[node] = trust[node]                   # device state is authoritative on this path
            tm._status[node] = status[node]
            tm._count[node] += 1
        if c.attack_detection_enabled:
            res = t.attack_detector.observe_stats(snap["act_stats"].numpy(), self.rank, step)
            if res.is_attack:
                t.metrics_collector.record_event("output_anomaly", node_id=self.rank, step=step)
        if "param_flags" in snap:
            # ZeRO-1: the parameter monitor ("Param Monitor", reference README.md:22) runs inside the parameter all-gather
            # of every step: a shard whose checksum differs from the one its owner published, or whose drift norm jumps,
            # marks the owner MODEL_POISONING in the same step
            pflags, pdrift = snap["param_flags"].tolist(), snap["param_drift"].tolist()
            for node in range(self.world):
                bad = bool(pflags[node])
                if bad:
                    t.parameter_monitor.report_mismatch(node, step)
                elif t.parameter_monitor.observe_drift(node, float(pdrift[node]), step)["poisoned"]:
                    bad = True
                if bad and tm.get_node_status(node) is not NodeStatus.COMPROMISED:
                    t.attack_history.append({"node_id": node, "timestamp": time.time(), "step": step,
                                             "attack_type": "model_poisoning",
                                             "parameter_stats": {"checksum_mismatch": bool(pflags[node]), "drift": float(pdrift[node])}})
                    tm.mark_compromised(node, "model_poisoning")
                    # the device-resident trust vector is authoritative on this path: demote the rank there too (every
                    # rank takes the same decision from the same flags, so the replicas stay identical)
                    self.engine.reducer.trust[node] = float(tm.get_trust_score(node))
                    self.engine.reducer.status[node] = 2
                    t.metrics_collector.record_event("attack", node_id=node, type="model_poisoning", step=step)
                    self.reassign(node)
                    from ..core.distributed_trainer import TrainingState
                    t.training_state = TrainingState.UNDER_ATTACK
        self._sec_s += time.perf_counter() - ts
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss, "step": step, "epoch": t.current_epoch, "trust_scores": dict(enumerate(trust)),
            "step_time": time.perf_counter() - t0, "samples": batch["input"].shape[0]})
        self._last_loss = loss
        return loss

    def _return_shards(self, step: int) -> None:
        """End of probation: a rank whose status is assignable again (TRUSTED / RECOVERING) and whose trust is back above
        the threshold takes its home shard back from whoever has been covering it (replicated decision, no traffic)."""
        t = self.t
        tm = t.trust_manager
        flagged = {a["node_id"] for a in t.attack_history if a["step"] == step}
        for node in range(self.world):
            if t.stage_map.stages_of(node) or node in flagged or not tm.can_assign_task(node) \
                    or tm.get_trust_score(node) < t.config.trust_threshold:
                continue
            cover = t.stage_map.owner[node]                      # shard i is rank i's home shard
            t.stage_map.owner[node] = node
            t.reassignment_history.append({"from_node": cover, "to_node": node, "timestamp": time.time(), "step": step,
                                           "stages": [node], "migration_time": 0.0, "kind": "shard_returned"})
            t.metrics_collector.record_event("reassignment", from_node=cover, to_node=node, step=step)
            t.defer_baseline_reset(cover)                        # both see a different batch from the next step on
            t.defer_baseline_reset(node)
            logger.info("node %d is trusted again (%.3f): data shard %d returned by node %d", node,
                        tm.get_trust_score(node), node, cover)

    def _check_replicas(self, step: int) -> None:
        """The parameter monitor of the replicated path ("Param Monitor", reference README.md:22; the reference has no
        code behind it).  Replicas apply the same aggregated gradient with the same optimizer, so their parameters
        are bit-identical by construction; before every forward each rank publishes the 64-bit checksum of its flat
        parameter vector (K9's checksum kernel, 16 bytes on the wire).  A rank that disagrees with the strict majority is
        flagged MODEL_POISONING before it computes anything, loses its data shard, and its replica is repaired by a
        broadcast from a majority rank.  Without a strict majority nobody is blamed; the most trusted rank's replica
        is taken as the reference."""
        from ..ops.integrity import checksum64
        t = self.t
        if t.attacker is not None:
            t.attacker.poison_parameters(self.rank, self.params, step)
        ck = checksum64(self.flat_params)
        table = comm.all_gather_vector(torch.tensor([ck & 0xFFFFFFFF, ck >> 32], dtype=torch.float64))
        keys = [(int(a), int(b)) for a, b in table.tolist()]
        votes: Dict[tuple, int] = {}
        for k in keys:
            votes[k] = votes.get(k, 0) + 1
        ref = max(votes, key=lambda k: (votes[k], -keys.index(k)))
        if votes[ref] == self.world:
            return
        trust = t.trust_manager.trust_vector()
        decisive = 2 * votes[ref] > self.world
        if not decisive:
            ref = keys[int(np.argmax(trust))]
            logger.warning("step %d: replicas disagree without a majority; resynchronising from the most trusted rank", step)
        src = max((r for r in range(self.world) if keys[r] == ref), key=lambda r: (trust[r], -r))
        if decisive:
            for node in (r for r in range(self.world) if keys[r] != ref):
                t.parameter_monitor.report_mismatch(node, step)
                if t.trust_manager.get_node_status(node).value == "compromised":
                    continue
                logger.error("Parameter poisoning detected on node %d", node)
                t.attack_history.append({"node_id": node, "timestamp": time.time(), "step": step,
                                         "attack_type": "model_poisoning",
                                         "parameter_stats": {"checksum_mismatch": True, "repaired_from": src}})
                t.metrics_collector.record_event("attack", node_id=node, type="model_poisoning", step=step)
                t.trust_manager.mark_compromised(node, "model_poisoning")
                self.reassign(node)
                from ..core.distributed_trainer import TrainingState
                t.training_state = TrainingState.UNDER_ATTACK
        comm.broadcast_tensor(self.flat_params, src)

    def _deviation(self, node: int, st) -> float:
        em, es = self.t.node_monitor.get_expected_mean(node), self.t.node_monitor.get_expected_std(node)
        if em is None or es is None:
            return 0.0
        return float(min(1.0, (abs(st[0] - em) / es + abs(st[1] - es) / es) / 2.0))

    def reassign(self, src: int) -> Optional[int]:
        """Hand the compromised rank's data shard(s) to the best trusted rank."""
        t = self.t
        moved = t.stage_map.stages_of(src)
        if not moved or not t.planner.allowed(src):
            return None
        loads = [float(len(t.stage_map.stages_of(n))) for n in range(self.world)]
        dst = t.planner.choose_target(src, loads)
        if dst is None:
            return None
        t.planner.note_attempt(src)
        t.stage_map.move_all(src, dst)
        t.reassignment_history.append({"from_node": src, "to_node": dst, "timestamp": time.time(),
                                       "step": t.global_step, "stages": moved, "migration_time": 0.0,
                                       "kind": "data_shard"})
        return dst

    @torch.no_grad()
    def validate(self, loader) -> float:
        model = self.t.model_partitions[0]
        if self.engine is not None:
            self.engine.export_to_module(model)       # the engine trains its own flat buffers: evaluate THOSE weights
        elif self.world > 1:
            # parameters are identical on every replica, BatchNorm running statistics are not (each rank saw its own
            # shard): evaluate with their mean over the replicas so that every rank reports the same number
            for b in model.buffers():
                if b.is_floating_point():
                    h = b.detach().cpu() if comm.is_gloo() and b.is_cuda else b.detach()
                    torch.distributed.all_reduce(h)
                    b.copy_(h.to(b.device) / self.world)
        model.eval()
        total, n = 0.0, 0
        for batch in loader:
            out = model(batch["input"].to(self.device))
            total += float(self.t.calculate_loss(out, batch["target"]).item())
            n += 1
        model.train()
        return total / max(1, n)

    def _consistency(self, cos, node: int, invalid: bool) -> float:
        """Gradient-consistency trust input: cosine to the median *relative to the typical rank* (noisy honest
        gradients all have a modest absolute cosine; only a rank far below the typical value is penalised)."""
        if invalid:
            return 0.0
        if self.world < 3:
            return 1.0
        fin = np.sort(np.asarray(cos, dtype=np.float64)[np.isfinite(cos)])
        med = float(fin[(fin.size - 1) // 2]) if fin.size else 1.0
        return float(np.clip(cos[node] / med, 0.0, 1.0)) if med > 0 else 1.0

    def comm_fraction(self) -> float:
        return self._sec_s / self._step_s if self._step_s > 0 else 0.0

    def cleanup(self) -> None:
        pass
