"""Inter-layer model parallelism across processes: ``parallelism: model`` (one micro-batch, the
reference's stage-after-stage flow) and ``parallelism: pipeline`` (micro-batched 1F1B).

The reference walks its stages in a python loop inside one process and sends nothing
(distributed_trainer.py:153-173).  Here each rank is a node owning one or more
stages; activations and activation-gradients move with NCCL / gloo p2p (row C1 of the SURVEY
work-list keeps PP traffic on NCCL by design); sends are non-blocking so they overlap the next
compute op.

Scheduling is data driven: a global topological order of ``(F|B, stage, micro-batch)`` ops is
generated once (GPipe order or a simulated 1F1B order) and each rank executes the ops of the
stages it owns.  Because the order is topological and sends never block, any stage->rank map is
deadlock free — which is what lets stages move between ranks (trust-driven reassignment) without
a restart.

Security is receiver-side: the rank that *consumes* an activation computes its statistics, so a
compromised sender cannot vouch for itself.  All ranks then all-gather the tiny per-stage
statistic vectors and per-node gradient norms and run the same detector / verifier / trust state
machine, i.e. decisions are replicated and deterministic.
"""
from __future__ import annotations

import logging
import time
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn

from ..core.partition import StageMap
from ..ops.reference_impl import NUM_STATS
from ..ops.stats import tensor_statistics
from . import comm
from ..utils.timing import nvtx_range

logger = logging.getLogger(__name__)

__all__ = ["PipelineStrategy", "make_schedule"]


def make_schedule(num_stages: int, micro_batches: int, kind: str = "1f1b") -> List[Tuple[str, int, int]]:
    """Global topological order of ops.  ``gpipe``: all forwards then all backwards.
    ``1f1b``: per-stage warm-up / steady 1F1B / cool-down sequences merged by simulation."""
    S, M = num_stages, micro_batches
    if kind == "gpipe" or M == 1:
        fwd = [("F", s, m) for m in range(M) for s in range(S)]
        bwd = [("B", s, m) for m in range(M) for s in reversed(range(S))]
        return fwd + bwd
    per_stage: List[List[Tuple[str, int]]] = []
    for s in range(S):
        warm = min(S - s - 1, M)
        ops: List[Tuple[str, int]] = [("F", m) for m in range(warm)]
        f, b = warm, 0
        while f < M:
            ops.append(("F", f)); f += 1
            ops.append(("B", b)); b += 1
        while b < M:
            ops.append(("B", b)); b += 1
        per_stage.append(ops)
    done_f, done_b = set(), set()
    ptr = [0] * S
    order: List[Tuple[str, int, int]] = []
    total = sum(len(o) for o in per_stage)
    while len(order) < total:
        progressed = False
        for s in range(S):
            if ptr[s] >= len(per_stage[s]):
                continue
            kind_, m = per_stage[s][ptr[s]]
            ready = (s == 0 or (s - 1, m) in done_f) if kind_ == "F" else \
                ((s, m) in done_f and (s == S - 1 or (s + 1, m) in done_b))
            if ready:
                order.append((kind_, s, m))
                (done_f if kind_ == "F" else done_b).add((s, m))
                ptr[s] += 1
                progressed = True
        if not progressed:
            raise RuntimeError("1F1B schedule simulation stalled")
    return order


class PipelineStrategy:
    def __init__(self, trainer):
        self.t = trainer
        self.rank, self.world = comm.rank(), comm.world_size()
        self.device = trainer.device
        self.stages: List[nn.Module] = []
        self._comm_s = 0.0
        self._step_s = 0.0
        self._grad_handoff: Dict[Tuple[int, int], torch.Tensor] = {}
        # message layouts both ends already know: key = (direction, stage boundary, micro-batch rows).  The first message
        # of a key travels with a shape/dtype header (read on the host: one synchronisation); afterwards only the payload
        # is sent and the receive needs no host round trip.  Cleared on every rank when the stage map changes.
        self._drift_flag = 0.0
        self._known_rx: Dict[Tuple[str, int, int], Tuple[Tuple[int, ...], torch.dtype]] = {}
        self._known_tx: set = set()

    # ------------------------------------------------------------------ setup
    def setup(self) -> None:
        t, c = self.t, self.t.config
        if c.num_nodes != self.world:
            logger.warning("num_nodes (%d) != world size (%d): using world size", c.num_nodes, self.world)
            c.num_nodes = self.world
            t.trust_manager.resize(self.world)
        # hybrid layout (BASELINE config 5): world = pp x dp; rank r is pipeline position r % pp of replica r // pp
        self.dp = max(1, int(c.dp_size)) if c.dp_size > 1 else 1
        assert self.world % self.dp == 0, "world size must be divisible by dp_size"
        self.pp = self.world // self.dp
        self.replica = self.rank // self.pp
        model = t._build_model()
        if hasattr(model, "untie_head") and self.pp > 1:
            model.untie_head()
        t.config.num_nodes, saved = self.pp, t.config.num_nodes
        parts = t.create_model_partitions()                         # pp stages per replica
        t.config.num_nodes = saved
        self.stages = [parts[i] for i in range(len(parts))]
        self.S = len(self.stages)
        # one global map: entry d*S + s is the rank that runs stage s of replica d
        t.stage_map = StageMap(self.dp * self.S, self.world)
        t.stage_map.owner = [d * self.pp + (s % self.pp) for d in range(self.dp) for s in range(self.S)]
        t.stage_costs = list(t.stage_costs) * self.dp
        for sid, st in enumerate(self.stages):
            t.model_partitions[sid] = st
            if self._owner(sid) == self.rank:
                self._adopt(sid)
        kind = "1f1b" if c.parallelism == "pipeline" else "gpipe"
        self.micro = max(1, c.micro_batches if c.parallelism == "pipeline" else 1)
        self.schedule = make_schedule(self.S, self.micro, kind)
        self._dp_groups: Dict[Tuple[int, ...], Any] = {}

    def _adopt(self, sid: int) -> None:
        st = self.stages[sid].to(self.device)
        self.t.parameter_monitor.reset(sid)              # a snapshot from an earlier tenure of this stage is stale
        params = [p for p in st.parameters() if p.requires_grad]
        if params and sid not in self.t.optimizers:
            self.t.optimizers[sid] = self.t._make_optimizer(params)
            self.t.schedulers[sid] = self.t._make_scheduler(self.t.optimizers[sid])

    def sync_ownership(self) -> None:
        """The stage map changed under us (a checkpoint written after a reassignment is being resumed): host exactly the
        stages the map assigns to this rank -- module on the device with an optimizer, everything else released."""
        for sid in range(self.S):
            if self._owner(sid) == self.rank:
                self._adopt(sid)
            elif sid in self.t.optimizers:
                self.t.optimizers.pop(sid)
                self.t.schedulers.pop(sid, None)
                self.stages[sid].to("cpu")
        self._known_rx.clear(); self._known_tx.clear()

    def _owner(self, sid: int, replica: Optional[int] = None) -> int:
        """Rank that runs stage ``sid`` of ``replica`` (default: this rank's replica)."""
        d = self.replica if replica is None else replica
        return self.t.stage_map.owner[d * self.S + sid]

    def _my_stages(self, rank: Optional[int] = None) -> List[int]:
        r = self.rank if rank is None else rank
        return [g % self.S for g in self.t.stage_map.stages_of(r)]

    def _dp_group(self, ranks: Tuple[int, ...]):
        """Process group of the ranks that hold the same stage in every replica (created collectively, cached)."""
        if ranks not in self._dp_groups:
            self._dp_groups[ranks] = dist.new_group(list(ranks))
        return self._dp_groups[ranks]

    def _sync_dp_grads(self, step: int) -> Dict[int, float]:
        """Trust-weighted aggregation of each stage's gradients across the data-parallel replicas (robust
        coefficients from the bucket norms exclude an exploding replica in the same step).  Returns the norm
        every participating rank contributed (own rank only)."""
        from ..ops.reference_impl import robust_coefficients, trust_weighted_reduce
        t = self.t
        tw_all = t.trust_manager.aggregation_weights()
        for sid in range(self.S):
            ranks = tuple(self._owner(sid, d) for d in range(self.dp))
            if len(set(ranks)) < len(ranks):
                continue
            group = self._dp_group(ranks)                            # every rank calls new_group in the same order
            if self.rank not in ranks:
                continue
            params = [p for p in self.stages[sid].parameters() if p.grad is not None]
            if not params:
                continue
            flat = torch.cat([p.grad.reshape(-1).float() for p in params])
            dev = torch.device("cpu") if comm.is_gloo() else flat.device
            flat_c = flat.to(dev)
            gathered = [torch.empty_like(flat_c) for _ in ranks]
            dist.all_gather(gathered, flat_c, group=group)
            norms = torch.stack([g.norm() for g in gathered])
            w = torch.tensor([tw_all[r] for r in ranks], dtype=torch.float32)
            w = w / w.sum() if float(w.sum()) > 0 else torch.full((len(ranks),), 1.0 / len(ranks))
            coeff = robust_coefficients(norms.cpu(), w)
            reduced, _ = trust_weighted_reduce(gathered, coeff.to(dev), 0.0, norms)
            reduced = reduced.to(flat.device)
            off = 0
            for p in params:
                n = p.numel()
                p.grad.copy_(reduced[off:off + n].view_as(p.grad).to(p.grad.dtype))
                off += n
        return {}

    def _send_msg(self, x: torch.Tensor, dst: int, tag: int, key: Tuple[str, int, int]) -> List:
        first = key not in self._known_tx
        self._known_tx.add(key)
        return comm.isend_tensor(x, dst, tag=tag, header=first)

    def _recv_msg(self, src: int, tag: int, key: Tuple[str, int, int]) -> torch.Tensor:
        like = self._known_rx.get(key)
        x = comm.recv_tensor(src, self.device, tag=tag, like=like)
        if like is None:
            self._known_rx[key] = (tuple(x.shape), x.dtype)
        return x

    # ------------------------------------------------------------- one step
    def train_step(self, batch: Dict[str, torch.Tensor]) -> float:
        t, c = self.t, self.t.config
        t0 = time.perf_counter()
        S, M = self.S, self.micro
        step = t.global_step
        xs = batch["input"].chunk(self.dp)[self.replica].chunk(M)
        ys = batch["target"].chunk(self.dp)[self.replica].chunk(M)
        # parameter monitor, exact half: nothing may have touched a hosted stage since its last optimizer step.  A stage that
        # fails is rolled back to the snapshot before this step's forward; the verdict travels with the decision vector.
        tampered = self._drift_flag
        self._drift_flag = 0.0
        if c.parameter_monitoring_enabled:
            for sid in self._my_stages():
                params = list(self.stages[sid].parameters())
                if t.attacker is not None:
                    t.attacker.poison_parameters(self.rank, params, step)
                if not t.parameter_monitor.unchanged(sid, params):
                    t.parameter_monitor.restore(sid, params)
                    tampered = 1.0
        acts: Dict[Tuple[int, int], Tuple[Optional[torch.Tensor], torch.Tensor]] = {}
        pending_sends: List = []
        stage_stats = torch.zeros(S, NUM_STATS, dtype=torch.float64)
        have_stats = torch.zeros(S, dtype=torch.float64)
        losses: List[torch.Tensor] = []
        stats_dev: Dict[int, torch.Tensor] = {}
        comm_s = 0.0
        for kind, s, m in self.schedule:
            if self._owner(s) != self.rank:
                continue
            if kind == "F":
                if s == 0:
                    x_in = xs[m].to(self.device)
                elif self._owner(s - 1) == self.rank:
                    x_in = acts[(s - 1, m)][1].detach()
                else:
                    tc = time.perf_counter()
                    with nvtx_range(f"pp.recv_act s{s} m{m}"):
                        x_in = self._recv_msg(self._owner(s - 1), 0, ("F", s - 1, xs[m].shape[0]))
                    comm_s += time.perf_counter() - tc
                if s > 0:
                    # receiver-side verification of the previous stage's output
                    if c.attack_detection_enabled and m == 0:
                        stats_dev[s - 1] = tensor_statistics(x_in)          # stays on the device until the schedule is issued
                        have_stats[s - 1] = 1.0
                    if x_in.is_floating_point():
                        x_in.requires_grad_(True)
                with nvtx_range(f"pp.fwd s{s} m{m}"), t._autocast():
                    out = self.stages[s](x_in)
                if t.attacker is not None:
                    out = t.attacker.poison_output(self.rank, out, step)
                acts[(s, m)] = (x_in if s > 0 else None, out)
                if s == S - 1:
                    if c.attack_detection_enabled and m == 0:
                        stats_dev[s] = tensor_statistics(out.detach())
                        have_stats[s] = 1.0
                    losses.append(t.calculate_loss(out, ys[m]) / M)
                elif self._owner(s + 1) != self.rank:
                    pending_sends += self._send_msg(out.detach(), self._owner(s + 1), 0, ("F", s, xs[m].shape[0]))
            else:
                x_in, out = acts.pop((s, m))
                if s == S - 1:
                    with nvtx_range(f"pp.bwd s{s} m{m}"):
                        losses[m].backward()
                else:
                    if self._owner(s + 1) == self.rank:
                        g = self._grad_handoff.pop((s + 1, m))
                    else:
                        tc = time.perf_counter()
                        with nvtx_range(f"pp.recv_grad s{s} m{m}"):
                            g = self._recv_msg(self._owner(s + 1), 1, ("B", s, xs[m].shape[0]))
                        comm_s += time.perf_counter() - tc
                    with nvtx_range(f"pp.bwd s{s} m{m}"):
                        out.backward(g)
                if s > 0 and x_in is not None and x_in.grad is not None:
                    if self._owner(s - 1) == self.rank:
                        self._grad_handoff[(s, m)] = x_in.grad
                    else:
                        pending_sends += self._send_msg(x_in.grad, self._owner(s - 1), 1, ("B", s - 1, xs[m].shape[0]))
        for w in pending_sends:
            if hasattr(w, "wait"):
                w.wait()
        self._grad_handoff.clear()
        if stats_dev:                                              # one device->host copy for all stage statistics
            order = sorted(stats_dev)
            stage_stats[order] = torch.stack([stats_dev[k].double() for k in order]).cpu()

        # ---- local gradients: fault injection + norm --------------------------------
        my_grads = [p.grad for sid in self._my_stages() for p in self.stages[sid].parameters() if p.grad is not None]
        if t.attacker is not None:
            t.attacker.poison_gradients(self.rank, my_grads, step)
        my_norm = t.gradient_verifier.total_norm(my_grads) if my_grads else 0.0
        if self.dp > 1:
            tc = time.perf_counter()
            self._sync_dp_grads(step)
            comm_s += time.perf_counter() - tc
        loss_local = float(sum(l.detach() for l in losses).item()) if losses else 0.0

        # ---- replicated decisions ---------------------------------------------------
        vec = torch.cat([stage_stats.reshape(-1), have_stats, torch.tensor([tampered, my_norm, loss_local,
                         1.0 if losses else 0.0], dtype=torch.float64)])
        tc = time.perf_counter()
        table = comm.all_gather_vector(vec).numpy()
        comm_s += time.perf_counter() - tc
        ss = table[:, :S * NUM_STATS].reshape(self.world, S, NUM_STATS)
        hv = table[:, S * NUM_STATS:S * NUM_STATS + S]
        norms = table[:, -3]
        loss = float(table[table[:, -1] > 0, -2].sum()) / self.dp  # mean over replicas
        tainted = self._judge_step(ss, hv, norms, step, tampered=table[:, -4])

        # ---- optimizer --------------------------------------------------------------
        for sid in list(t.optimizers):
            if self._owner(sid) != self.rank:
                continue
            opt = t.optimizers[sid]
            # a rank that failed verification drops its own step ONLY when it is the sole holder of the stage: with
            # dp > 1 every replica already holds the same trust-weighted aggregate (the coefficients excluded the bad
            # rank), and skipping the step on one replica would make the stage's copies diverge for good
            # a step in which a stage OUTPUT was flagged is dropped everywhere: every gradient of it descends from a
            # poisoned activation
            if tainted or (self.dp == 1 and self.rank in t._last_step_info["invalid_nodes"]):
                opt.zero_grad(set_to_none=True)
                continue
            if c.clip_norm > 0:
                nn.utils.clip_grad_norm_(self.stages[sid].parameters(), c.clip_norm)
            opt.step()
            opt.zero_grad(set_to_none=True)
        if c.parameter_monitoring_enabled:
            # drift half: the size of this update against the stage's own history; refreshes the snapshot the next step's
            # exact check compares with.  An abnormal update is rolled back and reported with the next decision vector.
            for sid in self._my_stages():
                params = list(self.stages[sid].parameters())
                if t.parameter_monitor.detect_parameter_poisoning(sid, params, step):
                    t.parameter_monitor.restore(sid, params)
                    self._drift_flag = 1.0
        self._comm_s += comm_s
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss, "step": step, "epoch": t.current_epoch,
            "trust_scores": {i: t.trust_manager.get_trust_score(i) for i in range(self.world)},
            "step_time": time.perf_counter() - t0, "samples": batch["input"].shape[0]})
        return loss

    def _judge_step(self, ss: np.ndarray, hv: np.ndarray, norms: np.ndarray, step: int,
                    tampered: Optional[np.ndarray] = None) -> bool:
        """The replicated decisions of one step, from the gathered table (identical on every rank): ``ss[src, s]`` = the 12
        statistics rank ``src`` measured on the output of stage ``s`` of its replica, ``hv`` says which entries exist,
        ``norms[node]`` = that rank's gradient norm.  Returns True when a stage output was flagged (the step is tainted).

        * Every node is judged before anything moves: offenders are recorded and demoted first, then their stages are
          reassigned, never onto a node flagged in the same step.
        * Stage outputs are judged in stage order, per replica, and only up to the first one flagged: everything computed
          from a poisoned activation -- the outputs of later stages, the loss, every gradient of that replica -- is
          tainted, so it is neither judged nor allowed into a baseline (an honest stage downstream of a poisoned one must
          not be blamed for the gradient explosion the poison causes)."""
        t, c, S = self.t, self.t.config, self.S
        out_stats: Dict[int, np.ndarray] = {}
        last_stage: Dict[int, int] = {}
        for s in range(S):
            for src in np.nonzero(hv[:, s] > 0)[0]:                 # reporter verified stage s of ITS replica
                node = self._owner(s, int(src) // self.pp)
                out_stats[node] = ss[src, s]                        # a node's last stage wins
                last_stage[node] = s
        t._last_step_info = {"out_stats": {n: (float(v[0]), float(v[1])) for n, v in out_stats.items()},
                             "invalid_nodes": set(), "grad_norms": {}}
        offenders: List[int] = []
        tainted_replicas: set = set()
        unjudged: set = set()
        if tampered is not None:
            # the parameter monitor of a rank found (and rolled back) parameters that changed outside the optimizer, or an
            # update far off the shard's drift history: MODEL_POISONING for that node; the step itself ran on clean weights
            for node in np.nonzero(np.asarray(tampered) > 0)[0]:
                node = int(node)
                t.attack_history.append({"node_id": node, "timestamp": time.time(), "step": step, "attack_type": "model_poisoning",
                                         "parameter_stats": {"restored_to_last_verified": True}})
                t.metrics_collector.record_event("attack", node_id=node, type="model_poisoning", step=step)
                t.trust_manager.mark_compromised(node, "model_poisoning")
                t.training_state = type(t.training_state)("under_attack")
                offenders.append(node)
        if c.attack_detection_enabled:
            for node in sorted(out_stats, key=lambda n: (n // self.pp, last_stage[n])):
                if node // self.pp in tainted_replicas:
                    unjudged.add(node)
                    continue
                res = t.attack_detector.observe_stats(out_stats[node], node, step)
                if res.is_attack:
                    if node not in offenders:
                        t.handle_detected_attack(node, None, result=res, reassign=False)
                        offenders.append(node)
                    tainted_replicas.add(node // self.pp)
        if c.gradient_verification_enabled:
            for node in range(self.world):
                if node in offenders or node // self.pp in tainted_replicas:
                    continue
                if t.stage_map.stages_of(node) and not t.gradient_verifier.verify_norm(float(norms[node]), node, step):
                    t._last_step_info["invalid_nodes"].add(node)
                    t.handle_gradient_attack(node, [], reassign=False)
                    offenders.append(node)
        for node in offenders:
            t.reassign_node_tasks(node)
        flagged = {a["node_id"] for a in t.attack_history if a["step"] == step}
        for node in range(self.world):
            if node // self.pp in tainted_replicas and node not in flagged:
                continue                                            # nothing this node produced in a tainted step is evidence
            st = t._last_step_info["out_stats"].get(node)
            dev = 1.0 if st is None else self._deviation(node, st)
            gc = 0.0 if not t.stage_map.stages_of(node) and node not in flagged else \
                (1.0 if node not in t._last_step_info["invalid_nodes"] else 0.0)
            t.trust_manager.update_trust_score(node, dev, gc)
            if st is not None and node not in flagged and node not in unjudged:
                t.node_monitor.record_output(node, *st)
        t._last_step_info["tainted_by"] = offenders[0] if tainted_replicas else None
        return bool(tainted_replicas)

    def _deviation(self, node: int, st) -> float:
        em, es = self.t.node_monitor.get_expected_mean(node), self.t.node_monitor.get_expected_std(node)
        if em is None or es is None:
            return 0.0
        return float(min(1.0, (abs(st[0] - em) / es + abs(st[1] - es) / es) / 2.0))

    # ---------------------------------------------------------- reassignment
    def reassign(self, src: int) -> Optional[int]:
        """Move every stage of ``src`` to the best trusted rank of the same pipeline replica: parameters are
        sent p2p; all ranks update the (replicated) stage map.  No process restarts."""
        t = self.t
        moved_global = t.stage_map.stages_of(src)
        if not moved_global or not t.planner.allowed(src):
            return None
        rep = src // self.pp
        loads = [t.stage_map.load(n, t.stage_costs) for n in range(self.world)]
        mean_load = (sum(loads) / max(1, sum(1 for l in loads if l > 0))) or 1.0
        cands = [n for n in t.trust_manager.get_trusted_nodes() if n != src and n // self.pp == rep]
        if not cands:
            logger.error("No trusted nodes available for reassignment in replica %d", rep)
            return None
        dst = max(cands, key=lambda n: (t.trust_manager.get_trust_score(n) - t.planner.load_weight * loads[n] / mean_load, -n))
        t.planner.note_attempt(src)
        t0 = time.perf_counter()
        moved = [g % self.S for g in moved_global]
        numel = sum(p.numel() for s2 in moved for p in self.stages[s2].parameters())
        for sid in moved:
            self._migrate(sid, src, dst)
        t.stage_map.move_all(src, dst)
        self._known_rx.clear(); self._known_tx.clear()             # new owners have not seen any header yet
        from ..core.reassignment import estimate_migration_time
        t.reassignment_history.append({
            "from_node": src, "to_node": dst, "timestamp": time.time(), "step": t.global_step, "stages": moved,
            "migration_time": estimate_migration_time(numel, 4, t.config.migration_model),
            "measured_time": time.perf_counter() - t0})
        t.defer_baseline_reset(dst)              # dst's baselines still judge what dst did in this step; reset before the next one
        if dst in t.node_configs:
            t.node_configs[dst].model_partition = "+".join(f"partition_{s2}" for s2 in self._my_stages(dst))
        return dst

    def _migrate(self, sid: int, src: int, dst: int) -> None:
        """Move stage ``sid`` with everything an identical continuation needs: parameters and buffers, the optimizer's
        per-parameter state (Adam moments / momentum buffers / step counts), its hyper-parameters as the scheduler left
        them, and the scheduler state (the reference checkpoints per-node optimizers, distributed_trainer.py:453-457, and
        "moves" a partition at :367-380 without any of it).  One flat fp32 payload with a checksum, plus a small pickled
        description of its layout."""
        import io
        import pickle
        st = self.stages[sid]
        if self.rank == src:
            chunks = [p.detach().reshape(-1).float() for p in st.state_dict().values() if p.is_floating_point()]
            dev = chunks[0].device if chunks else self.device
            opt = self.t.optimizers.pop(sid, None)
            sched = self.t.schedulers.pop(sid, None)
            meta = None
            if opt is not None:
                osd = opt.state_dict()
                layout, scalars = [], {}
                for idx in sorted(osd["state"]):
                    for k, v in sorted(osd["state"][idx].items()):
                        if torch.is_tensor(v):
                            layout.append((idx, k, tuple(v.shape), v.dtype))
                            chunks.append(v.detach().reshape(-1).float().to(dev))
                        else:
                            scalars[(idx, k)] = v
                meta = {"layout": layout, "scalars": scalars, "param_groups": osd["param_groups"],
                        "sched": sched.state_dict() if sched is not None else None}
            flat = torch.cat(chunks or [torch.zeros(1, device=dev)])
            cs = float(flat.double().sum().item())
            # GPU->GPU over NVLink under NCCL (the wire device is the stage's own device); host memory under gloo
            comm.send_tensor(torch.cat([flat, torch.tensor([cs], dtype=torch.float32, device=flat.device)]), dst, tag=2)
            blob = torch.frombuffer(bytearray(pickle.dumps(meta)), dtype=torch.uint8)
            comm.send_tensor(blob.to(dev), dst, tag=2)
            st.to("cpu")
            for p in st.parameters():
                p.grad = None
        elif self.rank == dst:
            buf = comm.recv_tensor(src, self.device, tag=2)
            meta = pickle.load(io.BytesIO(comm.recv_tensor(src, self.device, tag=2).cpu().numpy().tobytes()))
            flat = buf[:-1]
            if abs(float(flat.double().sum().item()) - float(buf[-1].item())) > 1e-3 * max(1.0, abs(float(buf[-1].item()))):
                raise RuntimeError(f"stage {sid}: parameter checksum mismatch after migration {src}->{dst}")
            off = 0
            sd = st.state_dict()
            for k, v in sd.items():
                if v.is_floating_point():
                    n = v.numel()
                    v.copy_(flat[off:off + n].reshape(v.shape).to(v.dtype))
                    off += n
            for p in st.parameters():
                p.grad = None
            self.t.optimizers.pop(sid, None)
            self.t.schedulers.pop(sid, None)
            self._adopt(sid)
            opt = self.t.optimizers.get(sid)
            if meta is not None and opt is not None:
                state: Dict[int, Dict] = {}
                for idx, k, shape, dtype in meta["layout"]:
                    n = 1
                    for d in shape:
                        n *= d
                    state.setdefault(idx, {})[k] = flat[off:off + n].reshape(shape).to(dtype).clone()
                    off += n
                for (idx, k), v in meta["scalars"].items():
                    state.setdefault(idx, {})[k] = v
                opt.load_state_dict({"state": state, "param_groups": meta["param_groups"]})
                if meta["sched"] is not None and self.t.schedulers.get(sid) is not None:
                    self.t.schedulers[sid].load_state_dict(meta["sched"])

    # ------------------------------------------------------------------ misc
    @torch.no_grad()
    def validate(self, loader) -> float:
        S = self.S
        total, n = 0.0, 0
        for batch in loader:
            x = None
            for s in range(S):
                if self._owner(s) != self.rank:
                    continue
                if s == 0:
                    x = batch["input"].chunk(self.dp)[self.replica].to(self.device)
                elif self._owner(s - 1) != self.rank:
                    x = comm.recv_tensor(self._owner(s - 1), self.device, tag=3)
                x = self.stages[s](x)
                if s < S - 1 and self._owner(s + 1) != self.rank:
                    comm.send_tensor(x, self._owner(s + 1), tag=3)
            loss = 0.0
            if self._owner(S - 1) == self.rank:
                loss = float(self.t.calculate_loss(x, batch["target"].chunk(self.dp)[self.replica]).item())
            vals = comm.all_gather_vector(torch.tensor([loss, 1.0 if self._owner(S - 1) == self.rank else 0.0]))
            total += float((vals[:, 0] * vals[:, 1]).sum() / max(1.0, float(vals[:, 1].sum())))
            n += 1
        return total / max(1, n)

    def comm_fraction(self) -> float:
        return self._comm_s / self._step_s if self._step_s > 0 else 0.0

    def cleanup(self) -> None:
        pass
