"""Tensor-parallel GPT-2 on the repo's own sm_100a kernels (BASELINE config 4: GPT-2 medium TP x8).

The reference has no tensor parallelism; the north star asks for fused GEMM -> reduce-scatter and
all-gather -> GEMM kernels on tcgen05/TMEM tiles fed by TMA.  Megatron layout with sequence parallelism: the
residual stream, both LayerNorms and every residual/bias add live on a **token shard** ``[M/W, C]``; attention and
the MLP run on **local heads / local hidden columns** over all ``M`` tokens.

Per block, forward (rank r):

    LayerNorm(shard) -> symmetric buffer ---barrier---> K8  AG(ln1) @ Wqkv_r^T + b    [M, 3C/W]   (one P2P pull, then the GEMM;
                                                                                        TDL_K8=tma: TMA reads of the peer tiles)
    attention on the C/W local heads (csrc/attention.cu)                               [M, C/W]
    K7  atty_r @ Wo_r^T -> TMA stores of partial tiles into the OWNER's slots ---barrier---> sum slots + bias + residual
    LayerNorm(shard) -> symmetric buffer ---barrier---> K8  AG(ln2) @ Wfc_r^T, GELU epilogue saving GELU'  [M, 4C/W]
    K7  h_r @ Wproj_r^T -> slots ---barrier---> sum slots + bias + residual                                [M/W, C]

Backward mirrors it: the dgrad of a row-parallel layer is a K8 over the gathered output gradient (weight read
MN-major, GELU' multiply and the bias gradient in the epilogue), the dgrad of a column-parallel layer is a K7;
wgrads are split-K tcgen05 GEMMs whose gathered operand arrives through a P2P pull (the K9 copy kernel) issued
beside the critical path; attention backward runs on the local heads.  One barrier (``tdl_exchange``) per fused
collective, none around the pulls (their buffers are not rewritten before a later barrier).  The whole step is
captured in one CUDA graph.

Replicated parameters (embeddings / tied head, LayerNorms, row-parallel biases) receive partial gradients from
every rank's token shard; they are aggregated by the trust-weighted all-reduce (K2) — which is where a rank that
injects poisoned gradients is caught — and rescaled by W (the aggregate is a weighted mean of partial sums).
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..models.gpt2 import GPT2Config
from ..ops import _lib, nn as K, tp_ops
from ..ops._lib import i32, ptr, stream_ptr
from ..ops.stats import tensor_statistics_device
from ..parallel.symm import SymmetricBuffer, SymmetricTrustReducer

__all__ = ["GPT2TPEngine"]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class GPT2TPEngine:
    def __init__(self, cfg: GPT2Config, batch_size: int, seq_len: int, lr: float = 5e-5, weight_decay: float = 0.01,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, device: Optional[torch.device] = None,
                 security: bool = True, use_graph: bool = True, seed: int = 0, clip_ratio: float = 2.0,
                 trust_threshold: float = 0.7):
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        W = self.world
        assert cfg.n_embd // cfg.n_head == 64 and cfg.n_head % W == 0, "local heads: n_head must divide by the TP degree (head_dim 64)"
        self.cfg = cfg
        self.B, self.T = batch_size, seq_len
        self.M = batch_size * seq_len
        assert self.M % (128 * W) == 0, "token shards must be multiples of 128 rows"
        assert batch_size % W == 0, "each rank's token shard must hold whole sequences (batch divisible by the TP degree)"
        self.Ms = self.M // W
        self.C, self.L, self.H = cfg.n_embd, cfg.n_layer, cfg.n_head
        self.Hl, self.Cl, self.Fl = self.H // W, self.C // W, 4 * self.C // W
        assert self.Cl % 32 == 0
        self.V, self.Vp = cfg.vocab_size, cfg.padded_vocab_size
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.security, self.use_graph = security, use_graph
        self.step_count = 0
        self._graph = None
        self._calls_per_step: Optional[int] = None
        self._build_params(seed, clip_ratio, trust_threshold)
        self._build_buffers()

    # ------------------------------------------------------------------ parameters
    def _specs(self):
        C, Cl, Fl = self.C, self.Cl, self.Fl
        sh_decay, sh_nodecay, rep_decay, rep_nodecay = [], [], [], []
        for l in range(self.L):
            p = f"h{l}."
            sh_decay += [(p + "qkv_w", (3 * Cl, C)), (p + "o_w", (C, Cl)), (p + "fc_w", (Fl, C)), (p + "proj_w", (C, Fl))]
            sh_nodecay += [(p + "qkv_b", (3 * Cl,)), (p + "fc_b", (Fl,))]
            rep_nodecay += [(p + "ln1_w", (C,)), (p + "ln1_b", (C,)), (p + "o_b", (C,)), (p + "ln2_w", (C,)), (p + "ln2_b", (C,)),
                            (p + "proj_b", (C,))]
        rep_decay += [("wte", (self.Vp, C)), ("wpe", (self.cfg.n_positions, C))]
        rep_nodecay += [("lnf_w", (C,)), ("lnf_b", (C,))]
        return sh_decay, sh_nodecay, rep_decay, rep_nodecay

    def _build_params(self, seed: int, clip_ratio: float, trust_threshold: float) -> None:
        dev = self.device
        sh_d, sh_n, rep_d, rep_n = self._specs()
        self.offsets: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        off = 0
        for group in (sh_d, sh_n):
            for name, shape in group:
                self.offsets[name] = (off, shape); off += math.prod(shape)
            if group is sh_d:
                self.sh_decay = off
        self.n_sharded = _round_up(off, 1024)
        off = self.n_sharded
        for group in (rep_d, rep_n):
            for name, shape in group:
                self.offsets[name] = (off, shape); off += math.prod(shape)
            if group is rep_d:
                self.rep_decay = off - self.n_sharded
        self.P = _round_up(off, 1024)
        self.n_rep = self.P - self.n_sharded
        self.p32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.p16 = torch.zeros(self.P, dtype=torch.bfloat16, device=dev)
        self.m = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.v = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.g32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.reducer = SymmetricTrustReducer(self.n_rep, torch.bfloat16, dev, clip_ratio=clip_ratio, trust_threshold=trust_threshold)
        self.bucket = self.reducer.full[: self.n_rep]
        self.hyper = torch.zeros(4, dtype=torch.float32, device=dev)
        self.attack_scale = torch.ones(1, dtype=torch.float32, device=dev)
        self._attack_scale_host = 1.0
        self.rep_scale = torch.full((1,), float(self.world), dtype=torch.float32, device=dev)
        # same full-model random stream on every rank; each rank keeps its slice (so TP == the unsharded model)
        g = torch.Generator(device="cpu").manual_seed(seed)
        C, L = self.C, self.L
        full: Dict[str, torch.Tensor] = {"wte": torch.randn(self.Vp * C, generator=g).mul_(0.02).view(self.Vp, C),
                                         "wpe": torch.randn(self.cfg.n_positions * C, generator=g).mul_(0.02).view(-1, C)}
        res_std = 0.02 / math.sqrt(2 * L)
        for l in range(L):
            full[f"h{l}.qkv_w"] = torch.randn(3 * C * C, generator=g).mul_(0.02).view(3 * C, C)
            full[f"h{l}.o_w"] = torch.randn(C * C, generator=g).mul_(res_std).view(C, C)
            full[f"h{l}.fc_w"] = torch.randn(4 * C * C, generator=g).mul_(0.02).view(4 * C, C)
            full[f"h{l}.proj_w"] = torch.randn(4 * C * C, generator=g).mul_(res_std).view(C, 4 * C)
        self.load_full_tensors(full, biases_default=True)

    def shard_of(self, name: str, t: torch.Tensor) -> torch.Tensor:
        """This rank's slice of a full-model tensor (column-parallel: output rows, row-parallel: input columns)."""
        r, Cl, Fl, C = self.rank, self.Cl, self.Fl, self.C
        key = name.split(".")[-1]
        if key == "qkv_w":
            return torch.cat([t[i * C + r * Cl:i * C + (r + 1) * Cl] for i in range(3)])
        if key == "qkv_b":
            return torch.cat([t[i * C + r * Cl:i * C + (r + 1) * Cl] for i in range(3)])
        if key == "o_w":
            return t[:, r * Cl:(r + 1) * Cl]
        if key == "fc_w" or key == "fc_b":
            return t[r * Fl:(r + 1) * Fl]
        if key == "proj_w":
            return t[:, r * Fl:(r + 1) * Fl]
        return t

    @torch.no_grad()
    def load_full_tensors(self, full: Dict[str, torch.Tensor], biases_default: bool = False) -> None:
        for name, (o, shape) in self.offsets.items():
            view = self.p32[o:o + math.prod(shape)].view(shape)
            if name in full:
                view.copy_(self.shard_of(name, full[name]).to(self.device, torch.float32))
            elif biases_default:
                view.fill_(1.0) if (name.endswith("_w") and len(shape) == 1) else view.zero_()
        if self.Vp > self.V:
            self.w32("wte")[self.V:].zero_()
        K.cast_f32_bf16(self.p32, self.p16)

    _MODULE_MAP = {"qkv_w": "attn.c_attn.weight", "qkv_b": "attn.c_attn.bias", "o_w": "attn.c_proj.weight",
                   "o_b": "attn.c_proj.bias", "fc_w": "mlp.c_fc.weight", "fc_b": "mlp.c_fc.bias",
                   "proj_w": "mlp.c_proj.weight", "proj_b": "mlp.c_proj.bias", "ln1_w": "ln_1.weight",
                   "ln1_b": "ln_1.bias", "ln2_w": "ln_2.weight", "ln2_b": "ln_2.bias"}

    def module_tensors(self, module) -> Dict[str, torch.Tensor]:
        out = {"wte": module.transformer.wte.weight, "wpe": module.transformer.wpe.weight,
               "lnf_w": module.transformer.ln_f.weight, "lnf_b": module.transformer.ln_f.bias}
        for l, blk in enumerate(module.transformer.h):
            sd = dict(blk.named_parameters())
            for short, long in self._MODULE_MAP.items():
                out[f"h{l}.{short}"] = sd[long]
        return out

    @classmethod
    def from_module(cls, module, batch_size: int, seq_len: int, **kw) -> "GPT2TPEngine":
        eng = cls(module.config, batch_size, seq_len, **kw)
        eng.load_full_tensors({k: v.detach() for k, v in eng.module_tensors(module).items()})
        return eng

    def w32(self, name): o, s = self.offsets[name]; return self.p32[o:o + math.prod(s)].view(s)
    def w16(self, name): o, s = self.offsets[name]; return self.p16[o:o + math.prod(s)].view(s)
    def grad32(self, name): o, s = self.offsets[name]; return self.g32[o:o + math.prod(s)].view(s)

    # ------------------------------------------------------------------ buffers
    def _build_buffers(self) -> None:
        dev, M, Ms, C, L, W = self.device, self.M, self.Ms, self.C, self.L, self.world
        Cl, Fl = self.Cl, self.Fl
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        n = Ms * C
        # one symmetric arena: per-layer LayerNorm outputs (K8 inputs, re-pulled by the wgrads), two gradient shards
        # (K8 inputs of the backward pass), the reduce-scatter staging slots
        self.arena = SymmetricBuffer((2 * L + 2) * n + W * n, torch.bfloat16, dev)
        a = self.arena.tensor
        self.sym_ln = [[a[(2 * l + j) * n:(2 * l + j + 1) * n].view(Ms, C) for j in range(2)] for l in range(L)]
        self.sym_dx = [a[(2 * L + j) * n:(2 * L + j + 1) * n].view(Ms, C) for j in range(2)]
        self.slots = a[(2 * L + 2) * n:(2 * L + 2) * n + W * n].view(W, Ms, C)
        self._off_ln = [[(2 * l + j) * n for j in range(2)] for l in range(L)]
        self._off_dx = [(2 * L + j) * n for j in range(2)]
        self._off_slots = (2 * L + 2) * n
        self.pad = SymmetricBuffer((int(_lib.lib().tdl_exchange_pad_bytes()) + 7) // 8, torch.int64, dev)
        self.epochs = torch.zeros(2, dtype=torch.int32, device=dev)
        self._nothing = torch.zeros(1, dtype=torch.float64, device=dev)
        self.idx = torch.zeros(self.B, self.T, dtype=torch.int64, device=dev)
        self.tgt = torch.zeros(self.B, self.T, dtype=torch.int64, device=dev)
        self.res = torch.empty(L + 1, Ms, C, **bf)
        self.x2 = torch.empty(L, Ms, C, **bf)
        self.qkv = torch.empty(L, M, 3 * Cl, **bf)
        self.atty = torch.empty(L, M, Cl, **bf)
        self.fcg = torch.empty(L, M, Fl, **bf)
        self.fch = torch.empty(L, M, Fl, **bf)
        self.ln_mean = torch.empty(2 * L + 1, Ms, **f32)
        self.ln_rstd = torch.empty(2 * L + 1, Ms, **f32)
        self.lse = torch.empty(L, self.B * self.Hl * self.T, **f32)
        self.lnf = torch.empty(Ms, C, **bf)
        self.logits = torch.empty(Ms, self.Vp, **bf)
        self.row_loss = torch.empty(Ms, **f32)
        self.loss_sum = torch.zeros(1, **f32)
        # gathered operands: the K8 pull lands every [M, C] activation / gradient here once; the forward copies are kept per
        # layer because the wgrads of the backward pass consume them again (no second trip over NVLink)
        self.k8_mode = os.environ.get("TDL_K8", "pull")            # pull | tma  (see _col_parallel)
        assert self.k8_mode in ("pull", "tma")
        self.k8_flags = self.k8_mode != "tma"                      # the gathered operand is kept locally
        self.ln_full = torch.empty(L, 2, M, C, **bf) if self.k8_flags else None
        self.full_a = torch.empty(M, C, **bf)
        self.full_b = torch.empty(M, C, **bf)
        self.dln_s = torch.empty(Ms, C, **bf)
        self.dfch = torch.empty(M, Fl, **bf)
        self.datty = torch.empty(M, Cl, **bf)
        self.dqkv = torch.empty(M, 3 * Cl, **bf)
        self.delta = torch.empty(self.B * self.Hl * self.T, **f32)
        self.act_stats = torch.zeros(12, dtype=torch.float64, device=dev)
        self.loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()
        self.idx_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.tgt_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.copy_stream = torch.cuda.Stream(device=dev)

    # ------------------------------------------------------------------ collectives
    def _peers(self, off: int) -> List[int]:
        return [p + off * 2 for p in self.arena.peer_ptrs]

    def barrier(self) -> None:
        if self.world > 1:
            _lib.call("tdl_exchange", self.pad.ptr_array(), i32(self.rank), i32(self.world), i32(0), ptr(self._nothing), i32(0),
                      ptr(self.epochs), i32(0), stream_ptr())

    def _col_parallel(self, off: int, w, out, full: Optional[torch.Tensor] = None, **kw) -> None:
        """K8: out [M, N_local] = AG(shards at arena offset ``off``) @ w.  Two forms (``TDL_K8``):

        * ``pull`` (default): one P2P pull brings the W-1 remote shards into the local buffer ``full`` (each remote byte
          crosses NVLink exactly once, at the copy kernel's ~0.5 TB/s), then the tcgen05 GEMM runs on local memory at full
          tile rate; ``full`` stays for the wgrad of the backward pass, so no operand is ever fetched twice.
        * ``tma``: the GEMM's TMA producer reads the peer shards directly (plain per-CTA loads, own shard first); the
          transfer is spread over the main loop, but an A tile crosses NVLink once per n-block (2-6x the bytes at GPT-2
          medium shapes) and the wgrads pull the operand again.

        A third form was built and measured and then removed: the pull as a small-footprint kernel on a second stream with
        a tile-arrival flag per 128-row block, the GEMM running at the same time and waiting per tile.  On B200 (2 GPUs,
        fc shape) it took 373 us against 68 us (pull) + 151 us (GEMM) back to back -- while a persistent tensor-core grid
        is resident the copy blocks only ran on the SMs held back for them -- and two launches of which one spins on the
        other have no co-scheduling guarantee on one GPU."""
        self.barrier()                                               # every rank's shard is written
        if full is None or self.k8_mode == "tma":
            tp_ops.allgather_gemm(self._peers(off), self.rank, self.Ms, w, out, **kw)
            if full is not None:
                tp_ops.gather_rows(self._peers(off), self.Ms * self.C, full.view(-1))
            return
        tp_ops.gather_rows(self._peers(off), self.Ms * self.C, full.view(-1))
        tp_ops.local_gemm(full, w, out, **kw)

    def _row_parallel(self, a, w, out_s, bias=None, residual=None, b_mn: bool = False) -> None:
        """K7 + slot sum: out_s [Ms, C] = RS(a @ w^T) (+ bias) (+ residual)."""
        tp_ops.gemm_scatter(a, w, self._peers(self._off_slots), self.rank, self.Ms, b_mn=b_mn)
        self.barrier()                                               # every partial tile has landed in its owner's slots
        tp_ops.reduce_slots(self.slots, self.world, out_s, residual=residual, bias=bias)

    def _gathered_ln(self, l: int, j: int) -> torch.Tensor:
        """The full [M, C] LayerNorm output of block ``l`` (j = 0: ln_1, 1: ln_2) for a wgrad: kept from the forward K8, or
        (TDL_K8=tma) pulled again from the per-layer symmetric shards."""
        if self.ln_full is not None:
            return self.ln_full[l, j]
        dst = self.full_a if j == 1 else self.full_b     # not in use at that point of the block's backward
        tmp = torch.empty_like(dst)
        tp_ops.gather_rows(self._peers(self._off_ln[l][j]), self.Ms * self.C, tmp.view(-1))
        return tmp

    def _lnf(self, l: int, j: int) -> Optional[torch.Tensor]:
        return self.ln_full[l, j] if self.ln_full is not None else None

    def _pull(self, off: int, dst: torch.Tensor, side: bool = True):
        """Gathered copy [M, C] of the shards at ``off`` for a wgrad.  Issued on the copy stream (beside the GEMMs of the
        critical path) after everything enqueued so far; returns the event the consumer waits for."""
        if not side:
            tp_ops.gather_rows(self._peers(off), self.Ms * self.C, dst.view(-1))
            return None
        main = torch.cuda.current_stream()
        ev0 = torch.cuda.Event(); ev0.record(main)
        self.copy_stream.wait_event(ev0)
        with torch.cuda.stream(self.copy_stream):
            tp_ops.gather_rows(self._peers(off), self.Ms * self.C, dst.view(-1))
            ev = torch.cuda.Event(); ev.record(self.copy_stream)
        return ev

    # ------------------------------------------------------------------ forward / backward
    def _forward(self) -> None:
        L, Ms, eps, r = self.L, self.Ms, self.cfg.layer_norm_eps, self.rank
        rows = slice(r * Ms, (r + 1) * Ms)
        idx_s = self.idx.view(-1)[rows]
        K.embedding_fwd(idx_s, self.w16("wte"), self.w16("wpe"), self.res[0], self.T)       # shards hold whole sequences
        for l in range(L):
            p = f"h{l}."
            K.layernorm_fwd(self.res[l], self.w32(p + "ln1_w"), self.w32(p + "ln1_b"), self.sym_ln[l][0],
                            self.ln_mean[2 * l], self.ln_rstd[2 * l], eps)
            # 256-wide tiles even when 3C/W is not a multiple of 256: every A tile crosses NVLink once per n-block
            self._col_parallel(self._off_ln[l][0], self.w16(p + "qkv_w"), self.qkv[l], full=self._lnf(l, 0),
                               bias=self.w32(p + "qkv_b"), block_n=256 if 3 * self.Cl > 128 else 0)
            K.attention_fwd(self.qkv[l], self.atty[l], self.lse[l], self.B, self.T, self.Hl)
            self._row_parallel(self.atty[l], self.w16(p + "o_w"), self.x2[l], bias=self.w32(p + "o_b"), residual=self.res[l])
            K.layernorm_fwd(self.x2[l], self.w32(p + "ln2_w"), self.w32(p + "ln2_b"), self.sym_ln[l][1],
                            self.ln_mean[2 * l + 1], self.ln_rstd[2 * l + 1], eps)
            self._col_parallel(self._off_ln[l][1], self.w16(p + "fc_w"), self.fcg[l], full=self._lnf(l, 1),
                               bias=self.w32(p + "fc_b"), epilogue=K.EPI_GELU_DG, aux_out=self.fch[l])
            self._row_parallel(self.fcg[l], self.w16(p + "proj_w"), self.res[l + 1], bias=self.w32(p + "proj_b"),
                               residual=self.x2[l])
        K.layernorm_fwd(self.res[L], self.w32("lnf_w"), self.w32("lnf_b"), self.lnf, self.ln_mean[2 * L], self.ln_rstd[2 * L], eps)
        K.linear_fwd(self.lnf, self.w16("wte"), self.logits)
        tgt_s = self.tgt.view(-1)[rows]
        K.xent_fwd_bwd(self.logits, tgt_s, self.row_loss, self.loss_sum, self.V, 1.0 / self.M)

    def _backward(self) -> None:
        L, Ms, r = self.L, self.Ms, self.rank
        K.linear_dgrad(self.logits, self.w16("wte"), self.dln_s)
        K.linear_wgrad(self.logits, self.lnf, self.grad32("wte"), split_k=1)
        cur = 0                                         # which symmetric gradient-shard buffer holds the running dx
        K.layernorm_bwd(self.dln_s, self.res[L], self.w32("lnf_w"), self.ln_mean[2 * L], self.ln_rstd[2 * L], None, self.sym_dx[cur],
                        self.grad32("lnf_w"), self.grad32("lnf_b"), dxsum=self.grad32(f"h{L - 1}.proj_b"))
        main = torch.cuda.current_stream()
        for l in reversed(range(L)):
            p = f"h{l}."
            # ---- MLP: d(fc out) = (AG(dx) @ Wproj_r) * gelu'; the gathered dx lands in full_a on the way and feeds the wgrad
            self._col_parallel(self._off_dx[cur], self.w16(p + "proj_w"), self.dfch, full=self.full_a, b_mn=True,
                               epilogue=K.EPI_MUL, aux_in=self.fch[l], colsum=self.grad32(p + "fc_b"))
            K.linear_wgrad(self.full_a, self.fcg[l], self.grad32(p + "proj_w"), split_k=0)
            K.linear_wgrad(self.dfch, self._gathered_ln(l, 1), self.grad32(p + "fc_w"), split_k=0)
            self._row_parallel(self.dfch, self.w16(p + "fc_w"), self.dln_s, b_mn=True)
            K.layernorm_bwd(self.dln_s, self.x2[l], self.w32(p + "ln2_w"), self.ln_mean[2 * l + 1], self.ln_rstd[2 * l + 1],
                            self.sym_dx[cur], self.sym_dx[cur ^ 1], self.grad32(p + "ln2_w"), self.grad32(p + "ln2_b"),
                            dxsum=self.grad32(p + "o_b"))
            cur ^= 1
            # ---- attention ----
            self._col_parallel(self._off_dx[cur], self.w16(p + "o_w"), self.datty, full=self.full_b, b_mn=True)
            K.linear_wgrad(self.full_b, self.atty[l], self.grad32(p + "o_w"), split_k=0)
            K.attention_bwd(self.qkv[l], self.atty[l], self.datty, self.lse[l], self.delta, self.dqkv, self.B, self.T, self.Hl,
                            dbias=self.grad32(p + "qkv_b"))
            K.linear_wgrad(self.dqkv, self._gathered_ln(l, 0), self.grad32(p + "qkv_w"), split_k=0)
            self._row_parallel(self.dqkv, self.w16(p + "qkv_w"), self.dln_s, b_mn=True)
            K.layernorm_bwd(self.dln_s, self.res[l], self.w32(p + "ln1_w"), self.ln_mean[2 * l], self.ln_rstd[2 * l],
                            self.sym_dx[cur], self.sym_dx[cur ^ 1], self.grad32(p + "ln1_w"), self.grad32(p + "ln1_b"),
                            dxsum=self.grad32(f"h{l - 1}.proj_b") if l > 0 else None)
            cur ^= 1
        rows = slice(r * Ms, (r + 1) * Ms)
        K.embedding_bwd(self.idx.view(-1)[rows], self.sym_dx[cur], self.grad32("wte"), self.grad32("wpe"), self.T)

    def _update(self) -> None:
        rd, b1, b2 = self.reducer, self.betas[0], self.betas[1]
        ns = self.n_sharded
        if self.security:
            tensor_statistics_device(self.res[self.L].view(-1), out=self.act_stats)
        K.adam_tick(self.hyper, self.lr, b1, b2)
        # sharded parameters: every rank owns its slice outright, fp32 gradients straight into AdamW
        K.adamw(self.p32[:ns], self.p16[:ns], self.g32[:ns], self.m[:ns], self.v[:ns], self.sh_decay, self.lr, b1, b2, self.eps,
                self.wd, 1, hyper=self.hyper)
        # replicated parameters: partial gradients -> trust-weighted aggregation (K1 + K2 + K3) -> x W -> AdamW
        sl = slice(ns, self.P)
        if self.world == 1:
            K.adamw(self.p32[sl], self.p16[sl], self.g32[sl], self.m[sl], self.v[sl], self.rep_decay, self.lr, b1, b2, self.eps,
                    self.wd, 1, hyper=self.hyper)
            return
        K.pack_norm(self.g32[sl], self.bucket, rd.sumsq, self.attack_scale)
        if self.security:
            rd.enqueue_bucket(0, self.n_rep)
            rd.enqueue_finalize(update_trust=True)
        else:
            rd.sumsq.zero_()
            rd.enqueue_sum_allreduce(0, self.n_rep)
        K.adamw(self.p32[sl], self.p16[sl], self.bucket, self.m[sl], self.v[sl], self.rep_decay, self.lr, b1, b2, self.eps,
                self.wd, 1, grad_scale=self.rep_scale, hyper=self.hyper)

    def _step_body(self) -> None:
        self.loss_sum.zero_()
        self.g32.zero_()
        self._forward()
        self._backward()
        self._update()

    # ------------------------------------------------------------------ public
    def run_step_device(self) -> None:
        if not self.use_graph:
            c0 = _lib.launch_count
            self._step_body()
            self._calls_per_step = _lib.launch_count - c0
        elif self._graph is None:
            if self.step_count == 0:
                self._step_body()
                torch.cuda.synchronize()
                self.step_count += 1
                return
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                g = torch.cuda.CUDAGraph()
                c0 = _lib.launch_count
                with torch.cuda.graph(g, stream=s):
                    self._step_body()
                self._calls_per_step = _lib.launch_count - c0
            torch.cuda.current_stream().wait_stream(s)
            self._graph = g
            g.replay()
        else:
            self._graph.replay()
        self.step_count += 1

    def train_step(self, idx: torch.Tensor, targets: torch.Tensor, attack_scale: float = 1.0) -> float:
        """Every rank passes the SAME global batch; returns this rank's share of the mean loss (sum over ranks = loss)."""
        if attack_scale != self._attack_scale_host:
            self.attack_scale.fill_(float(attack_scale)); self._attack_scale_host = float(attack_scale)
        if idx.device.type == "cpu":
            self.idx_host.copy_(idx.view(self.B, self.T)); self.tgt_host.copy_(targets.view(self.B, self.T))
            self.idx.copy_(self.idx_host, non_blocking=True); self.tgt.copy_(self.tgt_host, non_blocking=True)
        else:
            self.idx.copy_(idx.view(self.B, self.T)); self.tgt.copy_(targets.view(self.B, self.T))
        self.run_step_device()
        self.loss_host.copy_(self.loss_sum, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(self.loss_host.item()) / self.M

    def kernels_per_step(self) -> int:
        if self._calls_per_step is None:
            raise RuntimeError("run at least two steps first")
        return int(self._calls_per_step)

    def security_snapshot(self) -> Dict[str, torch.Tensor]:
        r = self.reducer
        return {"trust": r.trust.cpu(), "status": r.status.cpu(), "stats": r.stats.cpu(), "coeff": r.coeff.cpu(),
                "flags": r.flags.cpu(), "act_stats": self.act_stats.cpu()}

    def comm_bytes_per_step(self) -> Dict[str, float]:
        """NVLink bytes this rank moves per step in each direction (roofline denominators of the fused paths)."""
        W, L = self.world, self.L
        shard = self.Ms * self.C * 2
        ag = (W - 1) * shard                       # one all-gather / reduce-scatter of a [M, C] bf16 tensor
        return {"k8_allgather_gemm": 4 * L * ag, "k7_gemm_scatter": 4 * L * ag, "wgrad_pulls": 0 if self.k8_flags else 4 * L * ag,
                "replicated_grad_allreduce": 2 * (W - 1) / W * self.n_rep * 2}
