"""Tensor parallelism with sequence-parallel activations (BASELINE config 4: GPT-2 medium TP x8).

The reference has no tensor parallelism.  Megatron-style layout: between blocks the
activations are sharded by tokens (``[M/W, C]``; LayerNorm / residual run on the shard);

* **column-parallel** linears (QKV, MLP up) consume the all-gathered activation:
  ``AG(x_s) @ W_shard^T``.  On B200s that is kernel K8 (``tdl_allgather_gemm_bf16``): the GEMM's TMA
  producer pulls every A tile straight from the owning peer's symmetric buffer, so the gather is
  spread over the GEMM main loop.
* **row-parallel** linears (attention out, MLP down) produce partial sums that are reduce-scattered:
  kernel K7 (``tdl_gemm_scatter_bf16``) stores each partial tile from the GEMM epilogue into the owner
  rank's staging slot over NVLink; after the barrier ``tdl_reduce_slots`` sums the W slots (fused with
  the bias / residual add).

Backward uses the transposed pattern (dgrad of a column-parallel layer is a GEMM->reduce-scatter, dgrad of
a row-parallel layer an all-gather->GEMM).  On CPU / gloo the same autograd functions fall back to
``all_gather_into_tensor`` / ``reduce_scatter_tensor`` so the strategy is testable without GPUs.

Parameters that stay replicated (embeddings / tied head, LayerNorms, row-parallel biases) get their
gradients aggregated by the trust-weighted all-reduce — that is where a TP rank injecting poisoned
gradients is caught (coefficient 0, trust drop).
"""
from __future__ import annotations

import logging
import math
import time
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from ..core.partition import StageMap
from . import comm
from .trust_allreduce import TrustAllReduce

logger = logging.getLogger(__name__)

__all__ = ["TensorParallelStrategy", "ColumnParallelLinear", "RowParallelLinear", "TPGPT2"]


def _world() -> int:
    return comm.world_size()


def _all_gather_rows(x: torch.Tensor) -> torch.Tensor:
    W = _world()
    if W == 1:
        return x
    out = torch.empty((x.shape[0] * W,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x.contiguous())
    return out


def _reduce_scatter_rows(x: torch.Tensor) -> torch.Tensor:
    W = _world()
    if W == 1:
        return x
    out = torch.empty((x.shape[0] // W,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if comm.is_gloo():                                   # gloo has no reduce_scatter: all_reduce + slice
        full = x.contiguous().clone()
        dist.all_reduce(full)
        out.copy_(full.chunk(W)[comm.rank()])
    else:
        dist.reduce_scatter_tensor(out, x.contiguous())
    return out


class _FusedTP:
    """Symmetric staging buffers + barriers for the fused K7/K8 path (CUDA + NCCL only)."""

    def __init__(self, max_rows_shard: int, max_cols: int, device: torch.device):
        from .symm import SymmetricBuffer
        from ..ops import _lib
        self.W, self.rank = _world(), comm.rank()
        self.rows, self.cols = max_rows_shard, max_cols
        self.shard = SymmetricBuffer(max_rows_shard * max_cols, torch.bfloat16, device)          # AG source
        self.slots = SymmetricBuffer(self.W * max_rows_shard * max_cols, torch.bfloat16, device)  # RS staging
        pad_bytes = int(_lib.lib().tdl_exchange_pad_bytes())
        self.pad = SymmetricBuffer((pad_bytes + 7) // 8, torch.int64, device)
        self.epochs = torch.zeros(2, dtype=torch.int32, device=device)
        self.dummy = torch.zeros(1, dtype=torch.float64, device=device)

    def barrier(self) -> None:
        from ..ops import _lib
        from ..ops._lib import i32, ptr, stream_ptr
        _lib.call("tdl_exchange", self.pad.ptr_array(), i32(self.rank), i32(self.W), i32(0), ptr(self.dummy), i32(1),
                  ptr(self.epochs), i32(0), stream_ptr())

    def allgather_gemm(self, x_s: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
        from ..ops import tp_ops
        rows, K = x_s.shape
        self.barrier()                                              # peers finished reading the previous contents
        self.shard.tensor[: rows * K].view(rows, K).copy_(x_s)
        self.barrier()                                              # every shard is published
        out = torch.empty(rows * self.W, w.shape[0], dtype=torch.bfloat16, device=x_s.device)
        tp_ops.allgather_gemm(self.shard.peer_ptrs, self.rank, rows, w, out, bias=bias)
        return out

    def gemm_reduce_scatter(self, y: torch.Tensor, w: torch.Tensor, b_mn: bool = False) -> torch.Tensor:
        from ..ops import tp_ops
        M = y.shape[0]
        rows = M // self.W
        N = w.shape[1] if b_mn else w.shape[0]
        self.barrier()                                              # owners consumed their previous slots
        tp_ops.gemm_scatter(y, w, self.slots.peer_ptrs, self.rank, rows, b_mn=b_mn)
        self.barrier()                                              # all partial tiles have landed
        out = torch.empty(rows, N, dtype=torch.bfloat16, device=y.device)
        tp_ops.reduce_slots(self.slots.tensor, self.W, out)
        return out


_FUSED: Optional[_FusedTP] = None


def _fusable(x: torch.Tensor, w: torch.Tensor) -> bool:
    return (_FUSED is not None and x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
            and x.shape[0] % 128 == 0 and w.shape[0] % 32 == 0 and w.shape[1] % 64 == 0)


class _ColumnParallel(torch.autograd.Function):
    """y[M, N/W] = AG_rows(x_s) @ w^T + b ;  dx_s = RS_rows(dy @ w)."""

    @staticmethod
    def forward(ctx, x_s, w, b):
        ctx.save_for_backward(x_s, w)
        ctx.has_bias = b is not None
        if _fusable(x_s, w) and _FUSED.rows * _FUSED.cols >= x_s.numel():
            return _FUSED.allgather_gemm(x_s.contiguous(), w, b)
        return F.linear(_all_gather_rows(x_s), w, b)

    @staticmethod
    def backward(ctx, dy):
        x_s, w = ctx.saved_tensors
        dy = dy.contiguous()
        if _fusable(dy, w.t()) and dy.shape[0] // _world() * w.shape[1] <= _FUSED.rows * _FUSED.cols and w.shape[1] % 32 == 0:
            dx_s = _FUSED.gemm_reduce_scatter(dy, w, b_mn=True)                 # dy [M, N'] @ w [N', K]
        else:
            dx_s = _reduce_scatter_rows(dy @ w)
        x_full = _all_gather_rows(x_s)
        dw = dy.reshape(-1, dy.shape[-1]).t() @ x_full.reshape(-1, x_full.shape[-1])
        db = dy.reshape(-1, dy.shape[-1]).sum(0) if ctx.has_bias else None
        return dx_s, dw.to(w.dtype), db


class _RowParallel(torch.autograd.Function):
    """out_s[M/W, N] = RS_rows(y @ w^T) ;  dy = AG_rows(dout_s) @ w."""

    @staticmethod
    def forward(ctx, y, w):
        ctx.save_for_backward(y, w)
        if _fusable(y, w) and y.shape[0] // _world() * w.shape[0] <= _FUSED.rows * _FUSED.cols:
            return _FUSED.gemm_reduce_scatter(y.contiguous(), w)
        return _reduce_scatter_rows(F.linear(y, w))

    @staticmethod
    def backward(ctx, dout_s):
        y, w = ctx.saved_tensors
        dfull = _all_gather_rows(dout_s.contiguous())
        dy = dfull @ w
        dw = dfull.reshape(-1, dfull.shape[-1]).t() @ y.reshape(-1, y.shape[-1])
        return dy, dw.to(w.dtype)


class ColumnParallelLinear(nn.Module):
    """Holds rows ``[rank*N/W, (rank+1)*N/W)`` of the full ``[N, K]`` weight."""

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        super().__init__()
        self.weight = nn.Parameter(weight.detach().clone())
        self.bias = nn.Parameter(bias.detach().clone()) if bias is not None else None

    def forward(self, x_s):
        return _ColumnParallel.apply(x_s, self.weight, self.bias)


class RowParallelLinear(nn.Module):
    """Holds columns ``[rank*K/W, (rank+1)*K/W)`` of the full ``[N, K]`` weight; the bias is replicated and
    added once after the reduce-scatter."""

    def __init__(self, weight: torch.Tensor, bias: Optional[torch.Tensor]):
        super().__init__()
        self.weight = nn.Parameter(weight.detach().clone())
        self.bias = nn.Parameter(bias.detach().clone()) if bias is not None else None
        if self.bias is not None:
            self.bias.replicated = True

    def forward(self, y):
        out = _RowParallel.apply(y, self.weight)
        return out if self.bias is None else out + self.bias


class _TPBlock(nn.Module):
    def __init__(self, blk, rank: int, W: int, n_head: int):
        super().__init__()
        C = blk.ln_1.weight.numel()
        self.hl = n_head // W
        self.hd = C // n_head
        Cl = C // W
        self.ln_1, self.ln_2 = blk.ln_1, blk.ln_2
        for p in list(self.ln_1.parameters()) + list(self.ln_2.parameters()):
            p.replicated = True
        wq, wk, wv = blk.attn.c_attn.weight.detach().split(C, dim=0)
        bq, bk, bv = blk.attn.c_attn.bias.detach().split(C, dim=0)
        sl = slice(rank * Cl, (rank + 1) * Cl)
        self.c_attn = ColumnParallelLinear(torch.cat([wq[sl], wk[sl], wv[sl]]), torch.cat([bq[sl], bk[sl], bv[sl]]))
        self.attn_proj = RowParallelLinear(blk.attn.c_proj.weight.detach()[:, sl], blk.attn.c_proj.bias)
        F4 = blk.mlp.c_fc.weight.shape[0] // W
        fs = slice(rank * F4, (rank + 1) * F4)
        self.c_fc = ColumnParallelLinear(blk.mlp.c_fc.weight.detach()[fs], blk.mlp.c_fc.bias.detach()[fs])
        self.mlp_proj = RowParallelLinear(blk.mlp.c_proj.weight.detach()[:, fs], blk.mlp.c_proj.bias)

    def forward(self, x_s, B: int, T: int):
        qkv = self.c_attn(self.ln_1(x_s))                                  # [M, 3*C/W]
        Cl = qkv.shape[-1] // 3
        q, k, v = (t.view(B, T, self.hl, self.hd).transpose(1, 2) for t in qkv.split(Cl, dim=-1))
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B * T, Cl)
        x_s = x_s + self.attn_proj(y)
        h = F.gelu(self.c_fc(self.ln_2(x_s)), approximate="tanh")
        return x_s + self.mlp_proj(h)


class TPGPT2(nn.Module):
    """GPT-2 whose blocks are tensor-parallel over the default process group (token-sharded residual)."""

    def __init__(self, model, rank: int, W: int):
        super().__init__()
        cfg = model.config
        assert cfg.n_head % W == 0, "heads must divide the TP degree"
        self.cfg, self.rank, self.W = cfg, rank, W
        self.wte, self.wpe, self.ln_f = model.transformer.wte, model.transformer.wpe, model.transformer.ln_f
        for p in list(self.wte.parameters()) + list(self.wpe.parameters()) + list(self.ln_f.parameters()):
            p.replicated = True
        self.blocks = nn.ModuleList(_TPBlock(b, rank, W, cfg.n_head) for b in model.transformer.h)

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None):
        B, T = idx.shape
        M = B * T
        assert M % self.W == 0
        rows = M // self.W
        sl = slice(self.rank * rows, (self.rank + 1) * rows)
        pos = torch.arange(T, device=idx.device).repeat(B)
        x_s = self.wte(idx.reshape(-1)[sl]) + self.wpe(pos[sl])
        for blk in self.blocks:
            x_s = blk(x_s, B, T)
        logits = F.linear(self.ln_f(x_s), self.wte.weight)[..., : self.cfg.vocab_size]
        if targets is None:
            return logits
        loss_sum = F.cross_entropy(logits.float(), targets.reshape(-1)[sl], reduction="sum")
        return logits, loss_sum / M                                        # local share of the global mean


class TensorParallelStrategy:
    def __init__(self, trainer):
        self.t = trainer
        self.rank, self.world = comm.rank(), comm.world_size()
        self.device = trainer.device
        self._sec_s = self._step_s = 0.0
        self._last_loss = float("nan")

    def setup(self) -> None:
        global _FUSED
        t, c = self.t, self.t.config
        if c.num_nodes != self.world:
            c.num_nodes = self.world
            t.trust_manager.resize(self.world)
        base = t._build_model()
        assert hasattr(base, "transformer"), "tensor parallelism is implemented for the GPT-2 family"
        self.engine = None
        why = self._fused_blockers(base)
        if not why:
            # the whole tensor-parallel step on the repo's kernels (engine/gpt2_tp_engine.py): K7 / K8 for every fused
            # collective, tcgen05 wgrads, local-head attention, LayerNorm on the token shard, one CUDA graph
            self.engine = self._make_engine(base)
            self.model = base
            t.model_partitions[0] = base
            t.stage_map = StageMap(self.world, self.world)
            t.stage_costs = [1.0] * self.world
            return
        if c.engine == "fused":
            raise ValueError("engine: fused is not available for this tensor-parallel configuration: " + "; ".join(why))
        logger.info("tensor parallel: eager modules between the fused collectives (%s)", "; ".join(why))
        dtype = torch.bfloat16 if (c.dtype == "bf16" and self.device.type == "cuda") else torch.float32
        self.model = TPGPT2(base, self.rank, self.world).to(self.device, dtype)
        t.model_partitions[0] = self.model
        import os
        if (self.device.type == "cuda" and self.world > 1 and dist.get_backend() == "nccl" and dtype == torch.bfloat16
                and os.environ.get("TDL_TP_FUSED", "1") != "0"):
            T = c.seq_len or base.config.n_positions
            rows = c.batch_size * T // self.world
            _FUSED = _FusedTP(rows, 4 * base.config.n_embd, self.device)
        self.shard_params = [p for p in self.model.parameters() if not getattr(p, "replicated", False)]
        self.rep_params = [p for p in self.model.parameters() if getattr(p, "replicated", False)]
        n_rep = sum(p.numel() for p in self.rep_params)
        self.reducer = TrustAllReduce(n_rep, torch.float32 if dtype == torch.float32 else torch.bfloat16, self.device)
        t.optimizers[0] = t._make_optimizer(list(self.model.parameters()))
        t.schedulers[0] = t._make_scheduler(t.optimizers[0])
        t.stage_map = StageMap(self.world, self.world)
        t.stage_costs = [1.0] * self.world

    def _make_engine(self, base):
        """The tensor-parallel engine (one place, so that tests/test_fused_wrappers_cpu.py can run the host side of
        ``_engine_step`` over gloo with a stand-in)."""
        from ..engine.gpt2_tp_engine import GPT2TPEngine
        c = self.t.config
        return GPT2TPEngine.from_module(base, batch_size=c.batch_size, seq_len=c.seq_len or base.config.n_positions,
                                        lr=c.learning_rate, weight_decay=c.weight_decay, device=self.device,
                                        security=bool(c.attack_detection_enabled or c.gradient_verification_enabled),
                                        trust_threshold=c.trust_threshold)

    def _fused_blockers(self, base) -> List[str]:
        c, cfg = self.t.config, base.config
        T = c.seq_len or cfg.n_positions
        checks = [(self.device.type == "cuda" and dist.is_initialized() and dist.get_backend() == "nccl", "needs CUDA + NCCL"),
                  (c.dtype == "bf16", "dtype is not bf16"), (c.engine != "eager", "engine: eager"),
                  (str(c.optimizer).lower() == "adamw", f"optimizer {c.optimizer!r} (engine: adamw)"),
                  (str(c.lr_schedule).lower() in ("constant", "none", ""), f"lr_schedule {c.lr_schedule!r}"),
                  (not c.clip_norm, "clip_norm > 0"), (cfg.n_embd // cfg.n_head == 64, "head_dim != 64"),
                  (cfg.n_head % self.world == 0, "heads not divisible by the TP degree"),
                  (c.batch_size % self.world == 0, "batch not divisible by the TP degree"),
                  ((c.batch_size * T) % (128 * self.world) == 0 and T % 128 == 0, "token shard not a multiple of 128 rows")]
        return [m for ok, m in checks if not ok]

    def _engine_step(self, batch, t0) -> float:
        """Fused path: one graph replay per step; afterwards the host trust manager / histories are refreshed from the
        device-resident security state (a few hundred bytes)."""
        from ..core.trust_manager import NodeStatus
        t, c, eng = self.t, self.t.config, self.engine
        step = t.global_step
        if comm.is_short_batch(batch, c.batch_size):
            return self._last_loss
        scale = 1.0
        if t.attacker is not None and t.attacker._armed("gradient_poisoning", step, self.rank):
            i = t.attacker.config.intensity
            scale = -(1.0 + 4.0 * i) if t.attacker.config.gradient_mode == "sign_flip" else 1.0 + 20.0 * i
            t.attacker._record("gradient_poisoning", step, self.rank, mode=t.attacker.config.gradient_mode)
        eng.train_step(batch["input"], batch["target"], attack_scale=scale)
        lt = eng.loss_sum.double() / eng.M
        if self.world > 1:
            dist.all_reduce(lt)
        ts = time.perf_counter()
        snap = eng.security_snapshot()
        loss_val = float(lt.item())
        trust, status, flags = snap["trust"].tolist(), snap["status"].tolist(), snap["flags"].tolist()
        tm = t.trust_manager
        t._last_step_info = {"out_stats": {}, "invalid_nodes": set(), "grad_norms": {}}
        for node in range(self.world):
            if flags[node] & 1 and tm.get_node_status(node) is not NodeStatus.COMPROMISED:
                t._last_step_info["invalid_nodes"].add(node)
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
        self._sec_s += time.perf_counter() - ts
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss_val, "step": step, "epoch": t.current_epoch, "trust_scores": dict(enumerate(trust)),
            "step_time": time.perf_counter() - t0, "samples": batch["input"].shape[0]})
        self._last_loss = loss_val
        return loss_val

    def train_step(self, batch: Dict[str, torch.Tensor]) -> float:
        t, c = self.t, self.t.config
        t0 = time.perf_counter()
        step = t.global_step
        if self.engine is not None:
            return self._engine_step(batch, t0)
        idx, tgt = batch["input"].to(self.device), batch["target"].to(self.device)
        logits, loss = self.model(idx, tgt)
        loss.backward()
        # ---- replicated parameters: trust-weighted aggregation of their gradients -------------------
        ts = time.perf_counter()
        flat = torch.cat([p.grad.reshape(-1).to(self.reducer.buffer.dtype) if p.grad is not None
                          else torch.zeros(p.numel(), dtype=self.reducer.buffer.dtype, device=self.device) for p in self.rep_params])
        self.reducer.buffer[: flat.numel()].copy_(flat)
        if t.attacker is not None:
            t.attacker.poison_gradients(self.rank, [self.reducer.buffer], step)
        weights = torch.from_numpy(t.trust_manager.aggregation_weights()).float()
        stats, coeff = self.reducer.reduce_(weights)
        stats = stats.detach().cpu().double().numpy()
        # partial gradients sum over ranks: the weighted mean is rescaled by W (coefficients sum to one)
        scale = float(self.world)
        off = 0
        for p in self.rep_params:
            n = p.numel()
            p.grad = (self.reducer.buffer[off:off + n].view_as(p).to(p.dtype) * scale)
            off += n
        t._last_step_info = {"out_stats": {}, "invalid_nodes": set(), "grad_norms": {}}
        if c.gradient_verification_enabled:
            ok = t.gradient_verifier.verify_rank_stats(stats[:, 0], stats[:, 1], float(stats[0, 2]), step)
            for node in np.nonzero(~ok)[0]:
                node = int(node)
                if t.trust_manager.get_node_status(node).value != "compromised":
                    t._last_step_info["invalid_nodes"].add(node)
                    t.handle_gradient_attack(node, [])
        for node in range(self.world):
            gc = self._consistency(stats[:, 3], node, node in t._last_step_info["invalid_nodes"])
            t.trust_manager.update_trust_score(node, 0.0, gc)
        self._sec_s += time.perf_counter() - ts
        if c.clip_norm > 0:
            nn.utils.clip_grad_norm_(self.model.parameters(), c.clip_norm)
        t.optimizers[0].step()
        t.optimizers[0].zero_grad(set_to_none=True)
        loss_t = loss.detach().float().reshape(1)
        if self.world > 1:
            lt = loss_t.cpu() if comm.is_gloo() else loss_t
            dist.all_reduce(lt)
            loss_t = lt
        loss_val = float(loss_t.item())
        self._step_s += time.perf_counter() - t0
        t.metrics_collector.collect_batch_metrics({
            "loss": loss_val, "step": step, "epoch": t.current_epoch,
            "trust_scores": {i: t.trust_manager.get_trust_score(i) for i in range(self.world)},
            "step_time": time.perf_counter() - t0, "samples": idx.numel()})
        return loss_val

    def reassign(self, src: int) -> Optional[int]:
        """A tensor-parallel rank holds 1/W of every layer, so it cannot be dropped without re-sharding the whole model
        onto W - 1 ranks (head counts and hidden widths do not divide).  What the framework does instead -- and what the
        history entry says (``kind: tp_zero_weighted``) -- is zero-weight the rank: it keeps computing its slice, but its
        replicated-parameter gradients enter the aggregation with coefficient 0 from the same step on."""
        t = self.t
        if not t.planner.allowed(src):
            return None
        dst = t.planner.choose_target(src, [1.0] * self.world)
        if dst is None:
            return None
        t.planner.note_attempt(src)
        t.reassignment_history.append({"from_node": src, "to_node": dst, "timestamp": time.time(), "step": t.global_step,
                                       "stages": [], "migration_time": 0.0, "kind": "tp_zero_weighted"})
        return dst

    @torch.no_grad()
    def validate(self, loader) -> float:
        if self.engine is not None:
            return self._validate_engine(loader)
        total, n = 0.0, 0
        for batch in loader:
            _, loss = self.model(batch["input"].to(self.device), batch["target"].to(self.device))
            lt = loss.detach().float().reshape(1)
            if self.world > 1:
                lt = lt.cpu() if comm.is_gloo() else lt
                dist.all_reduce(lt)
            total += float(lt.item())
            n += 1
        return total / max(1, n)

    def _validate_engine(self, loader) -> float:
        """Forward-only through the engine (same kernels, same collectives), mean loss over the loader."""
        eng, total, n = self.engine, 0.0, 0
        for batch in loader:
            eng.idx.copy_(batch["input"].view(eng.B, eng.T)); eng.tgt.copy_(batch["target"].view(eng.B, eng.T))
            eng.loss_sum.zero_()
            eng._forward()
            lt = eng.loss_sum.double() / eng.M
            if self.world > 1:
                dist.all_reduce(lt)
            total += float(lt.item())
            n += 1
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
        global _FUSED
        _FUSED = None
