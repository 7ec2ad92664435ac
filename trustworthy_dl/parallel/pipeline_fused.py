"""GPT-2 pipeline / model parallelism on the repo's own kernels (BASELINE config 5: PP x4 x DP x2 with a mid-run
node compromise -> trust decay -> task reassignment; also ``parallelism: model`` for GPT-2 on GPUs).

The reference's only strategy is inter-layer model parallelism (stage split distributed_trainer.py:124-135, stage
loop :153-161, README.md:126) simulated in one process.  :class:`PipelineStrategy` (pipeline.py) runs arbitrary
``nn.Module`` stages; this subclass replaces the stage body, the transport and the gradient sync when the model is
GPT-2 in bf16 on NCCL:

* **stage body** = :class:`GPT2Engine` over a contiguous block range (tcgen05 GEMMs, attention, LayerNorm,
  cross-entropy, AdamW of csrc/): activations in and out as bf16 ``[micro_batch * T, C]`` buffers, one activation
  slot per in-flight micro-batch (1F1B: at most ``stages - s`` for stage ``s``);
* **transport** = NCCL p2p (SURVEY row C1 keeps it on NCCL) straight out of / into those buffers, no staging copy, no
  shape header, no host synchronisation: every (sender, receiver, direction) has its own communicator so forward
  and backward traffic never queue behind each other, receives are posted ahead (the next micro-batch's ``irecv`` is
  in flight while the current one computes) and consumed with a stream-level wait;
* **replica sync** = the trust-weighted all-reduce K2 over one world-symmetric arena: stage ``s`` always lives at
  offset ``s`` of whatever rank runs it, so the replicas of a stage find each other by pointer and a stage can
  change hands without a new rendezvous;
* **statistics** stay on the device (K5 on the received activation, bucket norms from K1) and cross PCIe once per
  step inside the replicated-decision gather;
* **reassignment** moves the fp32 master parameters, both Adam moments, the optimizer clock and the exchange epochs
  with the integrity-checked pull of K9 (checksum published by the sender, recomputed by the receiver), so training
  continues without a bump and without a restart.
"""
from __future__ import annotations

import logging
import os
import time
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ..core.partition import StageMap, balanced_partition
from ..ops import _lib
from ..ops._lib import i32, i64, ptr, stream_ptr
from ..ops.reference_impl import NUM_STATS
from ..ops.stats import tensor_statistics_device
from ..utils.timing import nvtx_range
from . import comm
from .pipeline import PipelineStrategy, make_schedule
from .symm import SymmetricBuffer

logger = logging.getLogger(__name__)

__all__ = ["FusedPipelineStrategy", "fused_pipeline_applicable", "split_layers", "can_post_early", "stage_slots"]

HEAD_COST = 2.7          # LM head + loss, fwd + bwd, in units of one transformer block (measured: 3 x 2.5 ms + 1.4 ms vs 3.3 ms)
EMBED_COST = 0.1


def fused_pipeline_applicable(trainer) -> Optional[str]:
    """None when the fused stage engine can run this configuration, else the reason it cannot."""
    c = trainer.config
    if trainer.device.type != "cuda" or not dist.is_initialized() or dist.get_backend() != "nccl":
        return "needs CUDA + NCCL"
    if not c.model_name.startswith("gpt2") or c.dtype != "bf16":
        return "fused stages exist for GPT-2 in bf16"
    if c.engine == "eager":
        return "engine: eager"
    if str(c.optimizer).lower() != "adamw" or str(c.lr_schedule).lower() not in ("constant", "none", ""):
        return "the stage engine runs AdamW at a constant learning rate"
    if c.clip_norm:
        return "clip_norm > 0"
    return None


def split_layers(n_layer: int, stages: int) -> List[Tuple[int, int]]:
    """Contiguous block ranges with the embedding on the first and ln_f + head on the last stage, balanced on
    measured cost (the head is worth ~2.7 blocks)."""
    costs = [1.0] * n_layer
    costs[0] += EMBED_COST
    costs[-1] += HEAD_COST
    b = balanced_partition(costs, stages)
    return list(zip(b[:-1], b[1:]))


def stage_slots(sid: int, num_stages: int, micro_batches: int, one_f_one_b: bool) -> int:
    """Activation slots of stage ``sid``: under 1F1B at most ``num_stages - sid`` micro-batches are in flight there,
    under GPipe order all of them."""
    return max(1, min(micro_batches, num_stages - sid)) if one_f_one_b else micro_batches


def can_post_early(cur: Tuple[str, int, int], nxt: Tuple[str, int, int], slots: int) -> bool:
    """May the receive of op ``nxt`` be posted BEFORE op ``cur`` is enqueued (so the transfer runs under ``cur``)?

    Not when ``cur`` still reads the buffer the receive writes: the same gradient buffer (a backward after a backward of
    the same stage with the same micro-batch parity) or the same activation slot (a forward that re-uses the slot of the
    micro-batch whose forward / backward runs now).  Ops of other stages never share buffers.  ``slots`` = activation
    slots of the stage of ``cur``.  Pure function: `tests/test_pipeline_plan.py` replays whole schedules against it."""
    kind, s, m = cur
    if nxt[1] != s:
        return True
    if nxt[0] == "B" and kind == "B" and nxt[2] % 2 == m % 2:
        return False
    if nxt[0] == "F" and nxt[2] % slots == m % slots:
        return False
    return True


class FusedPipelineStrategy(PipelineStrategy):
    # ------------------------------------------------------------------ setup
    def setup(self) -> None:
        t, c = self.t, self.t.config
        if c.num_nodes != self.world:
            c.num_nodes = self.world
            t.trust_manager.resize(self.world)
        self.dp = max(1, int(c.dp_size)) if c.dp_size > 1 else 1
        assert self.world % self.dp == 0, "world size must be divisible by dp_size"
        self.pp = self.world // self.dp
        self.replica = self.rank // self.pp
        model = t._build_model()
        self.full_model = model
        cfg = model.config
        self.gcfg = cfg
        if self.pp > 1:
            model.untie_head()
        # stages: one per pipeline rank, or `pp_virtual` per rank (interleaved, stage s starts on rank s % pp)
        self.S = S = min(self.pp * max(1, int(getattr(c, "pp_virtual", 1))), cfg.n_layer)
        self.ranges = split_layers(cfg.n_layer, S)
        kind = "1f1b" if c.parallelism == "pipeline" else "gpipe"
        self.micro = max(1, c.micro_batches if c.parallelism == "pipeline" else 1)
        self.schedule = make_schedule(S, self.micro, kind)
        self.T = c.seq_len or cfg.n_positions
        per_replica = c.batch_size // self.dp
        assert per_replica % self.micro == 0, "batch_size / dp_size must be divisible by micro_batches"
        self.mb = per_replica // self.micro
        t.stage_map = StageMap(self.dp * S, self.world)
        t.stage_map.owner = [d * self.pp + (s % self.pp) for d in range(self.dp) for s in range(S)]
        t.stage_costs = [float(b - a) + (EMBED_COST if s == 0 else 0.0) + (HEAD_COST if s == S - 1 else 0.0)
                         for s, (a, b) in enumerate(self.ranges)] * self.dp
        self._alloc_shared(cfg)
        self.n_buckets = 1 if self.dp == 1 else 2
        self.engines: Dict[int, Any] = {}
        # nn.Module view of each stage (public API: ``trainer.model_partitions``; numel for the migration model).  The live
        # weights are in the stage engines; :meth:`export_to_model` writes them back (checkpoints, validation on CPU).
        units = list(model.stage_units())                      # [embed, block 0 .. block L-1, head]
        self.stages = []
        for sid, (a, b) in enumerate(self.ranges):
            mods = ([units[0]] if sid == 0 else []) + units[1 + a:1 + b] + ([units[-1]] if sid == S - 1 else [])
            self.stages.append(torch.nn.Sequential(*mods))
            t.model_partitions[sid] = self.stages[sid]
            if self._owner(sid) == self.rank:
                self.engines[sid] = self._make_engine(sid, load=True)
        # ---- one communicator per FLOW = (sender rank, receiver rank, direction, stage boundary): a communicator carries the
        #      messages of one stage boundary one way only, in micro-batch order, so no send can wait behind an unmatched
        #      receive and two flows between the same pair of ranks can never be matched against each other.  Created for
        #      the flows the current stage map needs (every rank walks the same list); a reassignment adds its own.
        self._pair_groups: Dict[Tuple[int, int, int, int], Any] = {}
        self._last_loss = float("nan")
        self._trace = None
        if os.environ.get("TDL_PP_TRACE"):
            os.makedirs("gpurun_out", exist_ok=True)
            self._trace = open(f"gpurun_out/pp_trace_rank{self.rank}.txt", "w")
        self._ensure_groups()
        self._stats_buf = torch.zeros(S, NUM_STATS, dtype=torch.float64, device=self.device)
        self._have_buf = torch.zeros(S, dtype=torch.float64, device=self.device)
        self._dp_groups = {}

    def _alloc_shared(self, cfg) -> None:
        """One world-symmetric arena for the gradient buckets of every stage, one exchange pad per stage, one fp32 staging
        area for migrations; allocated once, sized for the largest stage."""
        from ..engine.gpt2_engine import SEG_ALIGN
        S = self.S
        C, Vp = cfg.n_embd, cfg.padded_vocab_size
        per_layer = 12 * C * C + 13 * C
        self.max_p = max((b - a) * per_layer + ((Vp + cfg.n_positions) * C if s == 0 else 0) + ((Vp + 2) * C if s == S - 1 else 0)
                         for s, (a, b) in enumerate(self.ranges)) + 8 * SEG_ALIGN
        self.arena = SymmetricBuffer(S * self.max_p, torch.bfloat16, self.device)
        pad_words = (int(_lib.lib().tdl_exchange_pad_bytes()) + 7) // 8
        self.pads = SymmetricBuffer(S * pad_words, torch.int64, self.device)
        self.mig = SymmetricBuffer(3 * self.max_p + 64, torch.float32, self.device, zero=False)

    def _sync_device(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize()

    def _tensor_stats(self, x: torch.Tensor, out: torch.Tensor) -> None:
        tensor_statistics_device(x, out=out)

    def _local_grad_sumsq(self, eng) -> None:
        """dp = 1: norm of the stage's fp32 gradients for the verifier (no bucket is packed before the verdict)."""
        _lib.call("tdl_sumsq", ptr(eng.g32), i32(0), i64(eng.P), ptr(eng.grad_sumsq), stream_ptr())

    def _ensure_groups(self) -> None:
        need = set()
        for d in range(self.dp):
            for sid in range(self.S - 1):
                a, b = self._owner(sid, d), self._owner(sid + 1, d)
                if a != b:
                    need.add((a, b, 0, sid)); need.add((b, a, 1, sid))
        fresh = []
        for key in sorted(need):
            if key not in self._pair_groups:
                self._pair_groups[key] = dist.new_group([key[0], key[1]])
                fresh.append(key)
        # First use of a communicator is a blocking host-side handshake that also allocates device buffers (an implicit
        # device synchronisation).  Done lazily inside a step it deadlocks as soon as receives are posted ahead: rank A
        # blocks in the handshake of its early irecv before it has enqueued the send that rank B's GPU is waiting for,
        # while B's handshake waits for B's GPU to drain.  So every new flow is exercised once here, with idle devices,
        # in one global order (each rank touches only its own flows, matched pairwise -> no cycle).
        if fresh:
            self._sync_device()
            dummy = torch.zeros(8, dtype=torch.bfloat16, device=self.device)
            for key in fresh:
                a, b = key[0], key[1]
                if self.rank == a:
                    dist.isend(dummy, b, group=self._pair_groups[key]).wait()
                elif self.rank == b:
                    dist.irecv(dummy, a, group=self._pair_groups[key]).wait()
            self._sync_device()

    def _dp_ranks(self, sid: int) -> List[int]:
        return [self._owner(sid, d) for d in range(self.dp)]

    def _slots(self, sid: int) -> int:
        return stage_slots(sid, self.S, self.micro, self.t.config.parallelism == "pipeline")

    def _make_engine(self, sid: int, load: bool, ranks: Optional[List[int]] = None):
        from ..engine.gpt2_engine import GPT2Engine
        c = self.t.config
        a, b = self.ranges[sid]
        eng = GPT2Engine(self.gcfg, batch_size=self.mb, seq_len=self.T, lr=c.learning_rate, weight_decay=c.weight_decay,
                         device=self.device, security=True, use_graph=False, layers=(a, b), has_embed=sid == 0,
                         has_head=sid == self.S - 1, tied_head=self.pp == 1, micro_slots=self._slots(sid),
                         micro_batches=self.micro, n_buckets=self.n_buckets, overlap=False,
                         dp_ranks=ranks if ranks is not None else self._dp_ranks(sid),
                         arena=self.arena, arena_offset=sid * self.max_p, pad=self.pads, pad_index=sid, init=False,
                         device_trust_update=False, trust_threshold=c.trust_threshold)
        if load:
            eng.load_from_module(self.full_model)
        return eng

    # ------------------------------------------------------------------ transport
    def _send(self, tensor: torch.Tensor, dst: int, direction: int, boundary: int):
        """``boundary`` = index of the lower stage of the stage boundary the message crosses."""
        self._log(f"isend dir{direction} b{boundary} -> {dst}")
        return dist.isend(tensor, dst, group=self._pair_groups[(self.rank, dst, direction, boundary)])

    def _recv(self, tensor: torch.Tensor, src: int, direction: int, boundary: int):
        self._log(f"irecv dir{direction} b{boundary} <- {src}")
        return dist.irecv(tensor, src, group=self._pair_groups[(src, self.rank, direction, boundary)])

    def _log(self, msg: str) -> None:
        if self._trace is not None:
            self._trace.write(f"{time.perf_counter():.4f} step {self.t.global_step} {msg}\n")
            self._trace.flush()

    # ------------------------------------------------------------- one step
    def train_step(self, batch: Dict[str, torch.Tensor]) -> float:
        t, c = self.t, self.t.config
        t0 = time.perf_counter()
        S, M = self.S, self.micro
        step = t.global_step
        if comm.is_short_batch(batch, c.batch_size):
            return self._last_loss
        xs = batch["input"].chunk(self.dp)[self.replica]
        ys = batch["target"].chunk(self.dp)[self.replica]
        mine = sorted(self.engines)
        scales: Dict[int, float] = {}
        my_ops = [(k, s, m) for (k, s, m) in self.schedule if s in self.engines]
        for sid in mine:
            eng = self.engines[sid]
            eng.begin_step()
            if eng.has_embed:                     # whole replica batch to the device once, one slot per micro-batch in flight
                self._idx_all = xs.to(self.device, non_blocking=True).view(M, self.mb, self.T)
            if eng.has_head:
                self._tgt_all = ys.to(self.device, non_blocking=True).view(M, self.mb, self.T)
            scale = 1.0
            if t.attacker is not None and t.attacker._armed("gradient_poisoning", step, self.rank):
                i = t.attacker.config.intensity
                scale = -(1.0 + 4.0 * i) if t.attacker.config.gradient_mode == "sign_flip" else 1.0 + 20.0 * i
                t.attacker._record("gradient_poisoning", step, self.rank, mode=t.attacker.config.gradient_mode)
            eng.attack_scale.fill_(scale)
            scales[sid] = scale
        self._stats_buf.zero_(); self._have_buf.zero_()
        # ---- receives are posted one op ahead of their consumer ----
        pending: Dict[Tuple[str, int, int], Any] = {}
        sends: List[Any] = []
        busy: Dict[Tuple[str, int, int], Any] = {}                  # send still reading (kind, stage, buffer index)

        def reuse(key):
            w0 = busy.pop(key, None)
            if w0 is not None:
                w0.wait()                                           # back-pressure: the buffer's previous send has drained

        def post_recv(op):
            kind, s, m = op
            eng = self.engines[s]
            if kind == "F" and s > 0 and self._owner(s - 1) != self.rank:
                pending[op] = self._recv(eng.res[m % eng.slots, 0], self._owner(s - 1), 0, s - 1)
            elif kind == "B" and s < S - 1 and self._owner(s + 1) != self.rank:
                pending[op] = self._recv(eng.dy_in[m % 2], self._owner(s + 1), 1, s)

        if my_ops:
            post_recv(my_ops[0])
        for i, op in enumerate(my_ops):
            kind, s, m = op
            eng = self.engines[s]
            slot = m % eng.slots
            # the receive buffer of the NEXT op is free unless this op still reads it: the same gradient buffer (B after B,
            # same parity) or the same activation slot (F re-using the slot whose backward runs now)
            nxt = my_ops[i + 1] if i + 1 < len(my_ops) else None
            early = nxt is not None and can_post_early(op, nxt, eng.slots)
            if early:
                post_recv(nxt)
            self._log(f"op {kind} s{s} m{m}")
            w = pending.pop(op, None)
            if w is not None:
                w.wait()                                            # stream-level: the compute stream waits for the transfer
            if kind == "F":
                reuse(("F", s, slot))
                if eng.has_embed:
                    eng.idx[slot].copy_(self._idx_all[m])
                elif self._owner(s - 1) == self.rank:
                    eng.res[slot, 0].copy_(self.engines[s - 1].res[m % self.engines[s - 1].slots, self.engines[s - 1].L])
                if s > 0 and c.attack_detection_enabled and m == 0:       # receiver-side check of the previous stage's output
                    self._tensor_stats(eng.res[slot, 0].view(-1), self._stats_buf[s - 1])
                    self._have_buf[s - 1] = 1.0
                with nvtx_range(f"pp.fwd s{s} m{m}"):
                    eng._forward(slot)
                out = eng.res[slot, eng.L]
                if t.attacker is not None:
                    o2 = t.attacker.poison_output(self.rank, out, step)
                    if o2 is not out:
                        out.copy_(o2)
                if eng.has_head:
                    eng.tgt[slot].copy_(self._tgt_all[m])
                    if c.attack_detection_enabled and m == 0:
                        self._tensor_stats(out.view(-1), self._stats_buf[s])
                        self._have_buf[s] = 1.0
                elif self._owner(s + 1) != self.rank:
                    busy[("F", s, slot)] = self._send(out, self._owner(s + 1), 0, s)
            else:
                k = m % 2
                reuse(("B", s, k))
                if not eng.has_head and self._owner(s + 1) == self.rank:
                    eng.dy_in[k].copy_(self.engines[s + 1].dx_out[k])
                with nvtx_range(f"pp.bwd s{s} m{m}"):
                    eng._backward(slot, k)
                if s > 0 and self._owner(s - 1) != self.rank:
                    busy[("B", s, k)] = self._send(eng.dx_out[k], self._owner(s - 1), 1, s - 1)
            if not early and nxt is not None:
                post_recv(nxt)
        for w in list(busy.values()) + sends:
            w.wait()
        # ---- gradient aggregation across the replicas of each stage (K1 + K2) + AdamW, or (dp = 1) decide first ----
        tw = t.trust_manager.aggregation_weights()
        norm_sq = torch.zeros(1, dtype=torch.float64, device=self.device)
        for sid in mine:
            eng = self.engines[sid]
            if self.dp > 1:
                ranks = self._dp_ranks(sid)
                w = torch.tensor([tw[r] for r in ranks], dtype=torch.float32)
                w = w / w.sum() * len(ranks) if float(w.sum()) > 0 else torch.ones(len(ranks))
                eng.reducer.trust.copy_(w.to(self.device), non_blocking=True)
                eng.reducer.status.zero_()
                with nvtx_range(f"pp.dp_sync s{sid}"):
                    eng.reduce_and_update()
            else:
                self._local_grad_sumsq(eng)
                if scales[sid] != 1.0:                              # the poisoned bucket is what this rank would ship
                    eng.grad_sumsq.mul_(scales[sid] ** 2)
            norm_sq += eng.grad_sumsq
        loss_dev = torch.zeros(2, dtype=torch.float64, device=self.device)
        if S - 1 in self.engines:
            loss_dev[0] = self.engines[S - 1].loss_sum[0].double() / (self.engines[S - 1].M * M)
            loss_dev[1] = 1.0
        # ---- replicated decisions: one small gather, one host sync per step ----
        vec = torch.cat([self._stats_buf.reshape(-1), self._have_buf, norm_sq.sqrt(), loss_dev])
        tc = time.perf_counter()
        table = comm.all_gather_vector(vec).numpy()
        comm_s = time.perf_counter() - tc
        ss = table[:, :S * NUM_STATS].reshape(self.world, S, NUM_STATS)
        hv = table[:, S * NUM_STATS:S * NUM_STATS + S]
        norms = table[:, -3]
        loss = float(table[table[:, -1] > 0, -2].sum()) / self.dp
        tainted = self._judge_step(ss, hv, norms, step)
        flagged = {a["node_id"] for a in t.attack_history if a["step"] == step}
        if self.dp == 1:
            # single holder of every stage: the (replicated) verdict gates the optimizer step of the poisoned rank
            # (a stage adopted from a flagged rank during this step arrives without gradients: its update is dropped)
            for sid in mine:
                if tainted or sid not in self.engines or self.rank in t._last_step_info["invalid_nodes"] or self.rank in flagged:
                    continue
                self.engines[sid].reduce_and_update()
        self._comm_s += comm_s
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss, "step": step, "epoch": t.current_epoch,
            "trust_scores": {i: t.trust_manager.get_trust_score(i) for i in range(self.world)},
            "step_time": time.perf_counter() - t0, "samples": batch["input"].shape[0]})
        self._last_loss = loss
        return loss

    # ---------------------------------------------------------- reassignment
    def reassign(self, src: int) -> Optional[int]:
        dst = super().reassign(src)
        if dst is not None:
            self._ensure_groups()
            # the replicas of every moved stage re-bind their peer pointers to the new owner
            for sid, eng in self.engines.items():
                if eng.world > 1:
                    eng.reducer.set_group(self._dp_ranks(sid))
        return dst

    def sync_ownership(self) -> None:
        """Resume of a checkpoint written after a reassignment: create the stage engines the loaded map assigns to this
        rank (their state follows through :meth:`load_checkpoint_state`), drop the others, open the flows of the new map."""
        for sid in range(self.S):
            mine = self._owner(sid) == self.rank
            if mine and sid not in self.engines:
                self.engines[sid] = self._make_engine(sid, load=False)
            elif not mine and sid in self.engines:
                del self.engines[sid]
        self._ensure_groups()
        for sid, eng in self.engines.items():
            if eng.world > 1:
                eng.reducer.set_group(self._dp_ranks(sid))

    def _migrate(self, sid: int, src: int, dst: int) -> None:
        """Move stage ``sid`` from rank ``src`` to ``dst``: parameters + Adam moments + optimizer clock + exchange epochs
        through the migration arena with the integrity-checked pull (K9).  Every rank takes part in the two barriers."""
        n = 0
        meta = torch.zeros(3, dtype=torch.int64, device=self.device)
        if self.rank == src:
            eng = self.engines.pop(sid)
            P = eng.P
            buf = self.mig.tensor
            buf[:P].copy_(eng.p32); buf[P:2 * P].copy_(eng.m); buf[2 * P:3 * P].copy_(eng.v)
            buf[3 * P:3 * P + 4].copy_(eng.hyper)
            buf[3 * P + 4:3 * P + 6].copy_(eng.reducer.epochs.float())
            buf[3 * P + 6] = float(eng.step_count)
            n = 3 * P + 8
            sums = torch.zeros(2, dtype=torch.int64, device=self.device)
            _lib.call("tdl_checksum64", ptr(buf), i64(n), ptr(sums), stream_ptr())
            meta[0], meta[1:] = n, sums & 0xFFFFFFFF
            del eng
        torch.cuda.synchronize()
        dist.broadcast(meta, src)                                   # size + the sender's checksum (also: staging is complete)
        n = int(meta[0].item())
        if self.rank == dst:
            ranks = [dst if r == src else r for r in self._dp_ranks(sid)]      # the stage map is updated after the move
            eng = self._make_engine(sid, load=False, ranks=ranks)
            P = eng.P
            assert n == 3 * P + 8, "stage layout mismatch between sender and receiver"
            import ctypes
            arr = (ctypes.c_void_p * 1)(self.mig.peer_ptrs[src])
            sums = torch.zeros(1, 2, dtype=torch.int64, device=self.device)
            drift = torch.zeros(1, dtype=torch.float64, device=self.device)
            local = self.mig.tensor
            _lib.call("tdl_param_gather", arr, i32(1), i32(0), ptr(local), None, i64(n), ptr(sums), ptr(drift), stream_ptr())
            got = (sums.view(-1) & 0xFFFFFFFF).cpu().tolist()
            want = meta[1:].cpu().tolist()
            if got != want:
                raise RuntimeError(f"stage {sid}: checksum mismatch after migration {src}->{dst}: {got} != {want}")
            eng.p32.copy_(local[:P]); eng.m.copy_(local[P:2 * P]); eng.v.copy_(local[2 * P:3 * P])
            eng.hyper.copy_(local[3 * P:3 * P + 4])
            eng.reducer.epochs.copy_(local[3 * P + 4:3 * P + 6].to(torch.int32))
            eng.step_count = int(local[3 * P + 6].item())
            eng.after_migration()
            self.engines[sid] = eng
        torch.cuda.synchronize()
        dist.barrier()                                              # the staging area may be reused

    # ------------------------------------------------------------------ misc
    def export_to_model(self) -> None:
        """Write the hosted stages' trained weights back into the nn.Module definition (checkpoints, validation)."""
        for eng in self.engines.values():
            eng.export_to_module(self.full_model)

    def checkpoint_state(self) -> Dict[str, Any]:
        """Per-rank file content: fp32 master weights, Adam moments and optimizer clock of every hosted stage."""
        return {"stages": {sid: {k: v.cpu() for k, v in eng.state_dict().items()} for sid, eng in self.engines.items()}}

    def load_checkpoint_state(self, sd: Dict[str, Any]) -> None:
        for sid, st in sd.get("stages", {}).items():
            if int(sid) in self.engines:
                eng = self.engines[int(sid)]
                eng.load_state_dict({k: (v.to(eng.device) if v.dim() else v) for k, v in st.items()})

    @torch.no_grad()
    def validate(self, loader) -> float:
        """Forward-only pass through the stage engines (same transport), mean loss over the loader."""
        S, total, n = self.S, 0.0, 0
        for batch in loader:
            xs = batch["input"].chunk(self.dp)[self.replica].to(self.device).view(self.micro, self.mb, self.T)
            ys = batch["target"].chunk(self.dp)[self.replica].to(self.device).view(self.micro, self.mb, self.T)
            loss_sum = torch.zeros(1, dtype=torch.float64, device=self.device)
            for m in range(self.micro):
                for s in range(S):
                    if s not in self.engines:
                        continue
                    eng = self.engines[s]
                    if eng.has_embed:
                        eng.idx[0].copy_(xs[m])
                    elif self._owner(s - 1) == self.rank:
                        eng.res[0, 0].copy_(self.engines[s - 1].res[0, self.engines[s - 1].L])
                    else:
                        self._recv(eng.res[0, 0], self._owner(s - 1), 0, s - 1).wait()
                    eng._forward(0)
                    if eng.has_head:
                        loss_sum += eng.head_loss_sum(0, ys[m])
                    elif self._owner(s + 1) != self.rank:
                        self._send(eng.res[0, eng.L], self._owner(s + 1), 0, s).wait()
            have = 1.0 if S - 1 in self.engines else 0.0
            vals = comm.all_gather_vector(torch.cat([loss_sum / (self.micro * self.mb * self.T),
                                                     torch.tensor([have], dtype=torch.float64, device=self.device)]))
            total += float((vals[:, 0] * vals[:, 1]).sum() / max(1.0, float(vals[:, 1].sum())))
            n += 1
        return total / max(1, n)

    def cleanup(self) -> None:
        self.engines.clear()
