"""Fused GPT-2 training engine for B200 (sm_100a).

This is the hot path behind ``DistributedTrainer`` when the model is GPT-2 on a GPU and behind
``bench.py``.  It replaces what the reference would run through torch eager / cuBLAS / ATen
(SURVEY 2.7, rows "stage forward", "loss", "backward", "optimizer", "gradient aggregation",
"per-stage stats") with an explicit forward/backward schedule over hand-written kernels:

* every GEMM (QKV, attention-out, MLP up/down, LM head; their dgrads and wgrads) is the tcgen05
  kernel in csrc/gemm_tcgen05.cu — bias, GELU (+saved derivative), residual add are epilogues;
  dgrad/wgrad consume the stored operands MN-major (no transposes); wgrads accumulate split-K into an
  fp32 main-grad buffer,
* attention is csrc/attention.cu (tcgen05 S=QK^T / PV / dK,dV / dQ, TMEM accumulators),
* LayerNorm fwd/bwd (bwd fused with the residual-gradient add), bias grads, embedding, fused
  cross-entropy and fused AdamW are csrc/elementwise.cu,
* security runs inside the step on the device: activation statistics (csrc/stats.cu), bucket
  pack + norm, trust-weighted reduction with fused median/cosine statistics over peer memory,
  device-side coefficient + trust state machine (csrc/trust_allreduce.cu).

The engine executes a **contiguous range of transformer blocks** (``layers=(l0, l1)``), optionally
with the embedding in front and ``ln_f`` + LM head + loss behind, over ``micro_slots`` activation
slots — so the same code is the whole model of the data-parallel headline path (one slot, whole step
in one CUDA graph) and one stage of the pipeline strategy (reference stage split
distributed_trainer.py:124-135, stage loop :153-161): activations enter and leave as bf16
``[micro_batch * T, C]`` buffers that NCCL p2p reads and writes in place.

**Gradient aggregation overlapped with backward** (README.md:31, docstring of
distributed_trainer.py:441-446): parameters are laid out in *segments* in the order their gradients
complete during backward (top block group first, embeddings last).  As soon as a segment's last
wgrad has been issued, a side stream packs it (K1), runs the trust-weighted reduction on that
sub-bucket (K2; per-bucket norms -> per-bucket coefficients) and applies AdamW to it, while the main stream
keeps back-propagating through lower blocks.  Measured: those kernels do not become resident beside the persistent
GEMM / attention CTAs, they take their turn at kernel boundaries, so most of the gain is the shorter exposed tail.  ``zero=1`` makes every rank keep and update
only its slice of each segment (ZeRO-1) and pull the others' bf16 parameters with the integrity-checked
parameter all-gather (K9: per-shard checksum vs the owner's published one + drift norm).
"""
from __future__ import annotations

import math
import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..utils.timing import nvtx_range
from ..models.gpt2 import GPT2Config
from ..ops import _lib, nn as K
from ..ops._lib import i32, i64, ptr, stream_ptr
from ..ops.stats import tensor_statistics_device
from ..parallel.symm import SymmetricBuffer, SymmetricTrustReducer

__all__ = ["GPT2Engine"]

SEG_ALIGN = 1024          # segment length granularity (elements): a multiple of 16-byte vectors x 8 ranks


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class _Segment:
    """A contiguous slice of the flat parameter buffer whose gradients complete together."""
    __slots__ = ("name", "off", "n", "n_decay", "tensors", "nd_off", "nd_n")

    def __init__(self, name: str):
        self.name, self.off, self.n, self.n_decay, self.tensors = name, 0, 0, 0, []
        self.nd_off = self.nd_n = 0            # this segment's 1-D tensors inside the fp32 view used by ZeRO-1


class GPT2Engine:
    def __init__(self, cfg: GPT2Config, batch_size: int, seq_len: int, lr: float = 5e-5, weight_decay: float = 0.01,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, device: Optional[torch.device] = None,
                 security: bool = True, use_graph: bool = True, seed: int = 0, clip_ratio: float = 2.0,
                 trust_threshold: float = 0.7, layers: Optional[Tuple[int, int]] = None, has_embed: bool = True,
                 has_head: bool = True, tied_head: Optional[bool] = None, micro_slots: int = 1, micro_batches: int = 1,
                 n_buckets: Optional[int] = None, overlap: Optional[bool] = None, zero: int = 0,
                 dp_ranks: Optional[Sequence[int]] = None, arena: Optional[SymmetricBuffer] = None, arena_offset: int = 0,
                 pad: Optional[SymmetricBuffer] = None, pad_index: int = 0, init: bool = True,
                 device_trust_update: bool = True):
        assert cfg.n_embd % cfg.n_head == 0 and cfg.n_embd // cfg.n_head == 64, "engine is specialised for head_dim 64"
        assert seq_len % 128 == 0 and seq_len <= cfg.n_positions
        self.cfg = cfg
        self.B, self.T = batch_size, seq_len
        self.M = batch_size * seq_len
        self.C, self.H = cfg.n_embd, cfg.n_head
        self.l0, self.l1 = layers if layers is not None else (0, cfg.n_layer)
        assert 0 <= self.l0 < self.l1 <= cfg.n_layer, "a stage needs at least one transformer block"
        self.L = self.l1 - self.l0                      # blocks executed by this engine
        self.has_embed, self.has_head = has_embed, has_head
        self.tied_head = (has_embed and has_head) if tied_head is None else (tied_head and has_embed and has_head)
        self.slots, self.micro_batches = int(micro_slots), int(micro_batches)
        self.head_in_backward = self.slots > 1 or self.micro_batches > 1     # one logits buffer: head runs right before its backward
        self.V, self.Vp = cfg.vocab_size, cfg.padded_vocab_size
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.security = security
        self.use_graph = use_graph and self.slots == 1 and self.micro_batches == 1
        self.zero = int(zero)
        self.device_trust_update = bool(device_trust_update)   # False: the host trust manager owns the trust vector (pipeline stages)
        self._init_params = bool(init)
        self.step_count = 0
        self._graph = None
        self._calls_per_step: Optional[int] = None
        if n_buckets is None:
            n_buckets = int(os.environ.get("TDL_BUCKETS", "4"))
        self.n_buckets = max(1, min(int(n_buckets), self.L))
        self.overlap = (os.environ.get("TDL_OVERLAP", "1") != "0") if overlap is None else bool(overlap)
        self._build_params(seed, trust_threshold, clip_ratio, dp_ranks, arena, arena_offset, pad, pad_index)
        self._build_activations()
        self.side = torch.cuda.Stream(device=self.device)

    # ------------------------------------------------------------------ parameters
    def _layer_specs(self, l: int):
        C = self.C
        p = f"h{l}."
        decay = [(p + "qkv_w", (3 * C, C)), (p + "o_w", (C, C)), (p + "fc_w", (4 * C, C)), (p + "proj_w", (C, 4 * C))]
        nodecay = [(p + "ln1_w", (C,)), (p + "ln1_b", (C,)), (p + "qkv_b", (3 * C,)), (p + "o_b", (C,)),
                   (p + "ln2_w", (C,)), (p + "ln2_b", (C,)), (p + "fc_b", (4 * C,)), (p + "proj_b", (C,))]
        return decay, nodecay

    def _plan_segments(self) -> List[_Segment]:
        """Segments in gradient-completion order: block groups from the top of the stage down (``ln_f`` and an untied
        head ride with the top group), then the embeddings (the tied ``wte`` receives its last contribution from the
        embedding backward)."""
        G, L = self.n_buckets, self.L
        bounds = [self.l1 - (L * g) // G for g in range(G + 1)]          # l1 = b0 > b1 > ... > bG = l0
        segs: List[_Segment] = []
        self._seg_of_layer: Dict[int, int] = {}
        for g in range(G):
            seg = _Segment(f"blocks[{bounds[g + 1]}:{bounds[g]})")
            decay, nodecay = [], []
            if g == 0 and self.has_head:
                if not self.tied_head:
                    decay.append(("lm_head", (self.Vp, self.C)))
                nodecay += [("lnf_w", (self.C,)), ("lnf_b", (self.C,))]
            for l in reversed(range(bounds[g + 1], bounds[g])):
                d, nd = self._layer_specs(l)
                decay += d; nodecay += nd
                self._seg_of_layer[l] = g
            seg.tensors = [(n, s, True) for n, s in decay] + [(n, s, False) for n, s in nodecay]
            segs.append(seg)
        if self.has_embed:
            seg = _Segment("embeddings")
            seg.tensors = [("wte", (self.Vp, self.C), True), ("wpe", (self.cfg.n_positions, self.C), True)]
            segs.append(seg)
        return segs

    def _build_params(self, seed: int, trust_threshold: float, clip_ratio: float, dp_ranks, arena, arena_offset, pad,
                      pad_index) -> None:
        dev = self.device
        self.segments = self._plan_segments()
        self.offsets: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        off = 0
        for seg in self.segments:
            seg.off = off
            for name, shape, decay in seg.tensors:
                n = math.prod(shape)
                assert n % 8 == 0
                self.offsets[name] = (off, shape)
                off += n
                if decay:
                    seg.n_decay = off - seg.off
            seg.nd_n = off - (seg.off + seg.n_decay)
            off = _round_up(off, SEG_ALIGN)
            seg.n = off - seg.off
        self.P = off
        nd = 0
        for seg in self.segments:
            seg.nd_off = nd
            nd += seg.nd_n
        self.p32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.m = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.v = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.g32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.reducer = SymmetricTrustReducer(self.P, torch.bfloat16, dev, clip_ratio=clip_ratio,
                                             trust_threshold=trust_threshold, ranks=dp_ranks, arena=arena,
                                             arena_offset=arena_offset, pad=pad, pad_index=pad_index)
        self.bucket = self.reducer.full[: self.P]
        self.world, self.rank = self.reducer.world, self.reducer.rank
        if self.zero:
            assert self.world & (self.world - 1) == 0, "ZeRO-1 slices need a power-of-two group"
            # the bf16 working copy lives in symmetric memory: peers pull each other's freshly updated slices
            self.p16_sym = SymmetricBuffer(self.P, torch.bfloat16, dev)
            self.p16 = self.p16_sym.tensor
            self.p16_prev = torch.zeros(self.P, dtype=torch.bfloat16, device=dev)
            G = len(self.segments)
            self.pg_sums = torch.zeros(G, self.world, 2, dtype=torch.int64, device=dev)     # receiver-side checksums
            self.pg_drift = torch.zeros(G, self.world, dtype=torch.float64, device=dev)     # ||theta - theta_prev||^2 per shard
            self.pg_own = torch.zeros(2, dtype=torch.int64, device=dev)                      # owner-side checksum (published)
            self.pg_pub = torch.zeros(G, self.world, 2, dtype=torch.int64, device=dev)       # what every owner published
            self.pg_flags = torch.zeros(self.world, dtype=torch.int32, device=dev)           # 1 = shard of that rank mismatched
            self.param_attack = torch.zeros(1, dtype=torch.float32, device=dev)              # fault injection: noise on my shard
            # ZeRO-1 publishes parameters in bf16, LayerNorm gains and biases included: every rank (the owner too, so the
            # replicas stay identical) feeds the fp32-consuming kernels from an fp32 view of the gathered bf16 values
            self.nd32 = torch.zeros(max(nd, 8), dtype=torch.float32, device=dev)
        else:
            self.p16 = torch.zeros(self.P, dtype=torch.bfloat16, device=dev)
        self.hyper = torch.zeros(4, dtype=torch.float32, device=dev)
        self.attack_scale = torch.ones(1, dtype=torch.float32, device=dev)      # fault-injection hook (1 = honest)
        self._attack_scale_host = 1.0
        self.grad_sumsq = torch.zeros(1, dtype=torch.float64, device=dev)       # this rank's squared gradient norm of the step
        # GPT-2 initialisation (same on every rank: replicas start identical); drawn per tensor in canonical model
        # order so a stage engine holds exactly the values the full-model engine would
        g = torch.Generator(device="cpu").manual_seed(seed)
        for name, shape in (self._canonical_order() if self._init_params else []):
            n = math.prod(shape)
            vals = None
            if len(shape) > 1:
                std = 0.02 / math.sqrt(2 * self.cfg.n_layer) if name.endswith(("o_w", "proj_w")) else 0.02
                vals = torch.randn(n, generator=g) * std
            targets = [name] if name in self.offsets else []
            if name == "wte" and "lm_head" in self.offsets:
                targets.append("lm_head")                  # an untied head starts as a copy of wte
            for t in targets:
                view = self.w32(t).view(-1)
                if vals is None:
                    view.fill_(1.0) if t.endswith("_w") else view.zero_()
                else:
                    view.copy_(vals.to(dev))
        for nm in ("wte", "lm_head"):
            if nm in self.offsets and self.Vp > self.V:
                self.w32(nm)[self.V:].zero_()
        self._refresh_working_copies()

    def _refresh_working_copies(self) -> None:
        """bf16 working parameters (and, under ZeRO-1, the fp32 view of the 1-D tensors) from the fp32 master."""
        K.cast_f32_bf16(self.p32, self.p16)
        if self.zero:
            self.p16_prev.copy_(self.p16)
            for seg in self.segments:
                self._refresh_nd(seg)

    def _refresh_nd(self, seg) -> None:
        if seg.nd_n:
            a = seg.off + seg.n_decay
            K.cast_bf16_f32(self.p16[a:a + seg.nd_n], self.nd32[seg.nd_off:seg.nd_off + seg.nd_n])

    def wf(self, name: str) -> torch.Tensor:
        """fp32 view of a 1-D tensor for the kernels that consume fp32 (LayerNorm gain / bias, GEMM bias)."""
        if not self.zero:
            return self.w32(name)
        o, shape = self.offsets[name]
        for seg in self.segments:
            if seg.off <= o < seg.off + seg.n:
                a = seg.nd_off + o - (seg.off + seg.n_decay)
                return self.nd32[a:a + math.prod(shape)].view(shape)
        raise KeyError(name)

    def _canonical_order(self):
        """(name, shape) of every tensor of the *whole* model in a fixed order (initialisation stream)."""
        C = self.C
        out = [("wte", (self.Vp, C)), ("wpe", (self.cfg.n_positions, C))]
        for l in range(self.cfg.n_layer):
            d, nd = self._layer_specs(l)
            out += d
        for l in range(self.cfg.n_layer):
            d, nd = self._layer_specs(l)
            out += nd
        out += [("lnf_w", (C,)), ("lnf_b", (C,))]
        return out

    def w32(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.p32[o:o + math.prod(shape)].view(shape)

    def w16(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.p16[o:o + math.prod(shape)].view(shape)

    def grad32(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.g32[o:o + math.prod(shape)].view(shape)

    @property
    def head_name(self) -> str:
        return "wte" if self.tied_head else "lm_head"

    # ------------------------------------------------------------------ module interop
    _MODULE_MAP = {"qkv_w": "attn.c_attn.weight", "qkv_b": "attn.c_attn.bias", "o_w": "attn.c_proj.weight",
                   "o_b": "attn.c_proj.bias", "fc_w": "mlp.c_fc.weight", "fc_b": "mlp.c_fc.bias",
                   "proj_w": "mlp.c_proj.weight", "proj_b": "mlp.c_proj.bias", "ln1_w": "ln_1.weight",
                   "ln1_b": "ln_1.bias", "ln2_w": "ln_2.weight", "ln2_b": "ln_2.bias"}

    def _module_tensors(self, module) -> Dict[str, torch.Tensor]:
        """Engine tensor name -> the matching parameter of a :class:`GPT2LMHeadModel` (only what this engine holds)."""
        out = {}
        if self.has_embed:
            out["wte"] = module.transformer.wte.weight
            out["wpe"] = module.transformer.wpe.weight
        if self.has_head:
            out["lnf_w"], out["lnf_b"] = module.transformer.ln_f.weight, module.transformer.ln_f.bias
            if not self.tied_head:
                hw = module.head.weight if getattr(module.head, "weight", None) is not None else module.transformer.wte.weight
                out["lm_head"] = hw
        for l in range(self.l0, self.l1):
            sd = dict(module.transformer.h[l].named_parameters())
            for short, long in self._MODULE_MAP.items():
                out[f"h{l}.{short}"] = sd[long]
        return out

    @classmethod
    def from_module(cls, module, batch_size: int, seq_len: int, **kw) -> "GPT2Engine":
        kw.pop("world", None); kw.pop("rank", None)
        eng = cls(module.config, batch_size, seq_len, **kw)
        eng.load_from_module(module)
        return eng

    @torch.no_grad()
    def load_from_module(self, module) -> None:
        for name, t in self._module_tensors(module).items():
            self.w32(name).copy_(t.detach().to(self.device, torch.float32))
        self._refresh_working_copies()

    @torch.no_grad()
    def sync_master(self) -> None:
        """ZeRO-1: every rank holds only its slice of the fp32 master weights and Adam moments; assemble the full buffers
        on every rank (NCCL all-gather, off the step path: checkpoints, export, validation on the nn.Module)."""
        if not self.zero or self.world == 1:
            return
        import torch.distributed as dist
        for seg in self.segments:
            per = seg.n // self.world
            lo = seg.off + self.reducer.rank * per
            for buf in (self.p32, self.m, self.v):
                dist.all_gather_into_tensor(buf[seg.off:seg.off + seg.n], buf[lo:lo + per].clone())

    @torch.no_grad()
    def export_to_module(self, module) -> None:
        self.sync_master()
        for name, t in self._module_tensors(module).items():
            t.copy_(self.w32(name).to(t.device, t.dtype))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        self.sync_master()
        return {"p32": self.p32.clone(), "m": self.m.clone(), "v": self.v.clone(), "hyper": self.hyper.clone(),
                "step": torch.tensor(self.step_count), "trust": self.reducer.trust.clone(),
                "status": self.reducer.status.clone()}

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self.p32.copy_(sd["p32"]); self.m.copy_(sd["m"]); self.v.copy_(sd["v"]); self.hyper.copy_(sd["hyper"])
        self.step_count = int(sd["step"])
        self.reducer.trust.copy_(sd["trust"]); self.reducer.status.copy_(sd["status"])
        self._refresh_working_copies()

    # ---- stage migration (trust-driven reassignment): everything a new owner needs to continue bit-identically ----
    def after_migration(self) -> None:
        """Called by the pipeline strategy once the stage's state (fp32 master parameters, both Adam moments, the optimizer
        clock, the cross-rank exchange epochs) has been copied into this engine's buffers."""
        self._refresh_working_copies()

    # ------------------------------------------------------------------ activations
    def _build_activations(self) -> None:
        dev, M, C, L, S = self.device, self.M, self.C, self.L, self.slots
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        if self.has_embed:
            self.idx = torch.zeros(S, self.B, self.T, dtype=torch.int64, device=dev)
        if self.has_head:
            self.tgt = torch.zeros(S, self.B, self.T, dtype=torch.int64, device=dev)
        self.res = torch.empty(S, L + 1, M, C, **bf)
        self.ln1 = torch.empty(S, L, M, C, **bf)
        self.qkv = torch.empty(S, L, M, 3 * C, **bf)
        self.atty = torch.empty(S, L, M, C, **bf)
        self.x2 = torch.empty(S, L, M, C, **bf)
        self.ln2 = torch.empty(S, L, M, C, **bf)
        self.fch = torch.empty(S, L, M, 4 * C, **bf)       # gelu'(fc pre-activation), consumed by the MLP backward
        self.fcg = torch.empty(S, L, M, 4 * C, **bf)
        self.ln_mean = torch.empty(S, 2 * L + 1, M, **f32)
        self.ln_rstd = torch.empty(S, 2 * L + 1, M, **f32)
        self.lse = torch.empty(S, L, self.B * self.H * self.T, **f32)
        if self.has_head:
            self.lnf = torch.empty(S, M, C, **bf)
            self.logits = torch.empty(M, self.Vp, **bf)
            self.row_loss = torch.empty(M, **f32)
        self.loss_sum = torch.zeros(1, **f32)
        # backward scratch
        self.dx_a = torch.empty(M, C, **bf)
        self.dx_b = torch.empty(M, C, **bf)
        self.dln = torch.empty(M, C, **bf)
        self.dqkv = torch.empty(M, 3 * C, **bf)
        self.datty = torch.empty(M, C, **bf)
        self.dfch = torch.empty(M, 4 * C, **bf)
        self.delta = torch.empty(self.B * self.H * self.T, **f32)
        if not self.has_head:
            self.dy_in = torch.zeros(2, M, C, **bf)         # gradient of the stage output (p2p receive buffers)
        if not self.has_embed:
            self.dx_out = torch.zeros(2, M, C, **bf)        # gradient of the stage input (p2p send buffers)
        self.act_stats = torch.zeros(12, dtype=torch.float64, device=dev)        # K5 on the stage output
        self.act_stats_in = torch.zeros(12, dtype=torch.float64, device=dev)     # K5 on the stage input (receiver-side check)
        self.act_stats_host = torch.zeros(12, dtype=torch.float64).pin_memory()
        self.sec_host = torch.zeros(self.world * 6, dtype=torch.float32).pin_memory()
        self.idx_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.tgt_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    def _split_k(self, n_out: int, k_in: int) -> int:
        """0 = let the kernel library choose the split that fills whole waves of the persistent grid."""
        return 0

    # ------------------------------------------------------------------ forward
    def _forward_blocks(self, s: int) -> None:
        L, eps = self.L, self.cfg.layer_norm_eps
        for j in range(L):
            p = f"h{self.l0 + j}."
            K.layernorm_fwd(self.res[s, j], self.wf(p + "ln1_w"), self.wf(p + "ln1_b"), self.ln1[s, j],
                            self.ln_mean[s, 2 * j], self.ln_rstd[s, 2 * j], eps)
            K.linear_fwd(self.ln1[s, j], self.w16(p + "qkv_w"), self.qkv[s, j], bias=self.wf(p + "qkv_b"))
            K.attention_fwd(self.qkv[s, j], self.atty[s, j], self.lse[s, j], self.B, self.T, self.H)
            K.linear_fwd(self.atty[s, j], self.w16(p + "o_w"), self.x2[s, j], bias=self.wf(p + "o_b"),
                         epilogue=K.EPI_RESIDUAL, aux_in=self.res[s, j])
            K.layernorm_fwd(self.x2[s, j], self.wf(p + "ln2_w"), self.wf(p + "ln2_b"), self.ln2[s, j],
                            self.ln_mean[s, 2 * j + 1], self.ln_rstd[s, 2 * j + 1], eps)
            K.linear_fwd(self.ln2[s, j], self.w16(p + "fc_w"), self.fcg[s, j], bias=self.wf(p + "fc_b"),
                         epilogue=K.EPI_GELU_DG, aux_out=self.fch[s, j])     # fch holds gelu'(pre): backward multiplies
            K.linear_fwd(self.fcg[s, j], self.w16(p + "proj_w"), self.res[s, j + 1], bias=self.wf(p + "proj_b"),
                         epilogue=K.EPI_RESIDUAL, aux_in=self.x2[s, j])
        if self.has_head:
            K.layernorm_fwd(self.res[s, L], self.wf("lnf_w"), self.wf("lnf_b"), self.lnf[s], self.ln_mean[s, 2 * L],
                            self.ln_rstd[s, 2 * L], eps)

    def _head_forward(self, s: int) -> None:
        """Logits + fused cross-entropy: leaves dlogits (scaled for the mean over the whole local batch) in ``logits``."""
        K.linear_fwd(self.lnf[s], self.w16(self.head_name), self.logits)
        K.xent_fwd_bwd(self.logits, self.tgt[s].view(-1), self.row_loss, self.loss_sum, self.V,
                       1.0 / (self.M * self.micro_batches))

    def _forward(self, s: int = 0) -> None:
        if self.has_embed:
            K.embedding_fwd(self.idx[s], self.w16("wte"), self.w16("wpe"), self.res[s, 0], self.T)
        self._forward_blocks(s)
        if self.has_head and not self.head_in_backward:
            self._head_forward(s)

    # ------------------------------------------------------------------ backward
    def _backward(self, s: int = 0, k: int = 0, on_segment_done=None) -> None:
        """Back-propagate micro-batch slot ``s``.  Without a head the output gradient is read from ``dy_in[k]``; without
        an embedding the input gradient is left in ``dx_out[k]``.  ``on_segment_done(i)`` fires once segment ``i``'s
        gradients are complete (every kernel that contributes to it has been enqueued)."""
        L, M, C = self.L, self.M, self.C
        done = on_segment_done or (lambda i: None)

        def scratch_other(cur):
            return self.dx_a if cur is not self.dx_a else self.dx_b

        top = f"h{self.l1 - 1}."
        if self.has_head:
            if self.head_in_backward:
                self._head_forward(s)
            dlogits = self.logits
            hn = self.head_name
            # LM head: dlnf = dlogits @ W ; dW += dlogits^T @ lnf
            K.linear_dgrad(dlogits, self.w16(hn), self.dln)
            K.linear_wgrad(dlogits, self.lnf[s], self.grad32(hn), split_k=1)
            dx = self.dx_a
            # every LayerNorm backward also emits the column sums of the dx it writes = the bias gradient of the linear
            # layer below it (MLP down-projection under ln_f / ln_1, attention output projection under ln_2)
            K.layernorm_bwd(self.dln, self.res[s, L], self.wf("lnf_w"), self.ln_mean[s, 2 * L], self.ln_rstd[s, 2 * L], None, dx,
                            self.grad32("lnf_w"), self.grad32("lnf_b"), dxsum=self.grad32(top + "proj_b"))
        else:
            dx = self.dy_in[k]
            K.bias_grad(dx, self.grad32(top + "proj_b"))              # the next stage's LayerNorm backward ran elsewhere
        for j in reversed(range(L)):
            l = self.l0 + j
            p = f"h{l}."
            # ---- MLP ----
            K.linear_wgrad(dx, self.fcg[s, j], self.grad32(p + "proj_w"), split_k=self._split_k(C, 4 * C))
            K.linear_dgrad(dx, self.w16(p + "proj_w"), self.dfch, epilogue=K.EPI_MUL, aux_in=self.fch[s, j],
                           colsum=self.grad32(p + "fc_b"))          # fc bias grad from the epilogue panels
            K.linear_wgrad(self.dfch, self.ln2[s, j], self.grad32(p + "fc_w"), split_k=self._split_k(4 * C, C))
            K.linear_dgrad(self.dfch, self.w16(p + "fc_w"), self.dln)
            nxt = scratch_other(dx)
            K.layernorm_bwd(self.dln, self.x2[s, j], self.wf(p + "ln2_w"), self.ln_mean[s, 2 * j + 1], self.ln_rstd[s, 2 * j + 1],
                            dx, nxt, self.grad32(p + "ln2_w"), self.grad32(p + "ln2_b"), dxsum=self.grad32(p + "o_b"))
            dx = nxt
            # ---- attention ----
            K.linear_wgrad(dx, self.atty[s, j], self.grad32(p + "o_w"), split_k=self._split_k(C, C))
            K.linear_dgrad(dx, self.w16(p + "o_w"), self.datty)
            K.attention_bwd(self.qkv[s, j], self.atty[s, j], self.datty, self.lse[s, j], self.delta, self.dqkv, self.B, self.T,
                            self.H, dbias=self.grad32(p + "qkv_b"))         # qkv bias grad from the accumulator write-out
            K.linear_wgrad(self.dqkv, self.ln1[s, j], self.grad32(p + "qkv_w"), split_k=self._split_k(3 * C, C))
            K.linear_dgrad(self.dqkv, self.w16(p + "qkv_w"), self.dln)
            nxt = self.dx_out[k] if (j == 0 and not self.has_embed) else scratch_other(dx)
            K.layernorm_bwd(self.dln, self.res[s, j], self.wf(p + "ln1_w"), self.ln_mean[s, 2 * j], self.ln_rstd[s, 2 * j],
                            dx, nxt, self.grad32(p + "ln1_w"), self.grad32(p + "ln1_b"),
                            dxsum=self.grad32(f"h{l - 1}.proj_b") if j > 0 else None)
            dx = nxt
            g = self._seg_of_layer[l]
            if j == 0 or self._seg_of_layer[l - 1] != g:
                done(g)
        if self.has_embed:
            K.embedding_bwd(self.idx[s], dx, self.grad32("wte"), self.grad32("wpe"), self.T)
            done(len(self.segments) - 1)

    # ------------------------------------------------------------------ aggregation + optimizer, one segment
    def _reduce_segment(self, i: int, overlapped: bool) -> None:
        """K1 -> K2 (or K4 with security off) -> AdamW on segment ``i``; everything is enqueued on the current stream."""
        seg, r = self.segments[i], self.reducer
        sl = slice(seg.off, seg.off + seg.n)
        b1, b2 = self.betas
        if self.world == 1:
            # nothing to aggregate: AdamW reads the fp32 main gradients directly (no bf16 rounding of the optimizer input);
            # the norm still feeds the verifier / trust update, the median of one rank is the rank itself
            if self.security:
                _lib.call("tdl_sumsq", ptr(self.g32[sl]), i32(0), i64(seg.n), ptr(r.sumsq), stream_ptr())
                self.grad_sumsq.add_(r.sumsq)
                r.partials.add_(r.sumsq)                     # dot(g, median) = |median|^2 = |g|^2
                r.acc_sq.add_(r.sumsq)
                r.sumsq.zero_()
            K.adamw(self.p32[sl], self.p16[sl], self.g32[sl], self.m[sl], self.v[sl], seg.n_decay, self.lr, b1, b2, self.eps,
                    self.wd, 1, grad_scale=self.attack_scale, hyper=self.hyper)
            return
        K.pack_norm(self.g32[sl], self.bucket[sl], r.sumsq, self.attack_scale)           # K1 (+ fault-injection scale)
        self.grad_sumsq.add_(r.sumsq)
        if self.security:
            r.enqueue_bucket(seg.off, seg.n, overlapped=overlapped, owner_only=bool(self.zero))      # K2
        else:
            r.sumsq.zero_()
            r.enqueue_sum_allreduce(seg.off, seg.n)                                                  # K4
        if not self.zero:
            K.adamw(self.p32[sl], self.p16[sl], self.bucket[sl], self.m[sl], self.v[sl], seg.n_decay, self.lr, b1, b2,
                    self.eps, self.wd, 1, hyper=self.hyper)
            return
        self._zero_update(i)

    def _zero_update(self, i: int) -> None:
        """ZeRO-1: AdamW on this rank's slice of segment ``i``, publish its checksum, pull every slice with K9."""
        seg, r, W = self.segments[i], self.reducer, self.world
        per = seg.n // W
        lo = seg.off + r.rank * per
        own = slice(lo, lo + per)
        b1, b2 = self.betas
        nd = min(max(seg.n_decay - r.rank * per, 0), per)
        K.adamw(self.p32[own], self.p16[own], self.bucket[own], self.m[own], self.v[own], nd, self.lr, b1, b2, self.eps,
                self.wd, 1, hyper=self.hyper)
        # fault-injection hook of the parameter-poisoning attack: corrupt my published shard AFTER its checksum ...
        self.pg_own.zero_()
        _lib.call("tdl_checksum64", ptr(self.p16[own]), i64(per // 2), ptr(self.pg_own), stream_ptr())
        _lib.call("tdl_poison_if", ptr(self.p16[own]), i64(per), ptr(self.param_attack), i64(1234 + i), stream_ptr())
        # ... publish (2 words through the exchange pad; doubles as the "slices are written" barrier) and gather
        _lib.call("tdl_exchange", r._pad_ptrs, i32(r.rank), i32(W), i32(1), ptr(self.pg_own), i32(2), ptr(r.epochs), i32(0),
                  stream_ptr())
        _lib.call("tdl_exchange_read", r._my_pad, i32(W), i32(1), i32(2), ptr(self.pg_pub[i]), stream_ptr())
        # peer k's freshly updated slice sits at offset k*per of the segment in peer k's copy of the bf16 parameters
        import ctypes
        peers = r.ranks if self.p16_sym.world > 1 else [0] * W
        arr = (ctypes.c_void_p * W)(*[self.p16_sym.peer_ptrs[peers[k]] + (seg.off + k * per) * 2 for k in range(W)])
        _lib.call("tdl_param_gather", arr, i32(W), i32(1), ptr(self.p16[seg.off:seg.off + seg.n]),
                  ptr(self.p16_prev[seg.off:seg.off + seg.n]), i64(per), ptr(self.pg_sums[i]), ptr(self.pg_drift[i]),
                  stream_ptr())
        _lib.call("tdl_integrity_compare", ptr(self.pg_sums[i]), ptr(self.pg_pub[i]), i32(W), ptr(self.pg_flags), stream_ptr())
        self._refresh_nd(seg)
        r.barrier(0)                                  # nobody overwrites a slice that a peer is still pulling

    # ------------------------------------------------------------------ step pieces
    def begin_step(self) -> None:
        self.loss_sum.zero_()
        self.g32.zero_()
        self.grad_sumsq.zero_()
        K.adam_tick(self.hyper, self.lr, self.betas[0], self.betas[1])
        if self.zero:
            self.pg_sums.zero_(); self.pg_drift.zero_(); self.pg_flags.zero_()

    def reduce_and_update(self, overlapped_already: bool = False) -> None:
        """Close the step: aggregate + update every segment (unless the backward pass already did, overlapped), then
        the security finalisation (cosines, flags, trust EMA + state machine) once for the whole step."""
        r = self.reducer
        if not overlapped_already:
            for i in range(len(self.segments)):
                self._reduce_segment(i, overlapped=False)
        if self.security:
            r.enqueue_finalize(update_trust=self.device_trust_update)
        elif self.world > 1:
            r.partials.zero_(); r.acc_sq.zero_(); r.acc_bad.zero_()

    def _step_body(self) -> None:
        self.begin_step()
        with nvtx_range("engine.forward"):
            self._forward(0)
        if self.security:
            tensor_statistics_device(self.res[0, self.L].view(-1), out=self.act_stats)       # K5 on the last residual
        overlap = self.overlap and len(self.segments) > 1
        with nvtx_range("engine.backward(+overlapped aggregation)"):
            if overlap:
                main = torch.cuda.current_stream()
                last = len(self.segments) - 1

                def seg_done(i):
                    ev = torch.cuda.Event()
                    ev.record(main)
                    self.side.wait_event(ev)
                    with torch.cuda.stream(self.side):
                        self._reduce_segment(i, overlapped=False)

                self._backward(0, 0, seg_done)
                main.wait_stream(self.side)
            else:
                self._backward(0, 0)
        with nvtx_range("engine.finalize"):
            self.reduce_and_update(overlapped_already=overlap)

    # ------------------------------------------------------------------ public step
    def run_step_device(self) -> None:
        """Enqueue one full training step on whatever is in ``self.idx`` / ``self.tgt`` (no host sync)."""
        if not self.use_graph:
            c0 = _lib.launch_count
            self._step_body()
            self._calls_per_step = _lib.launch_count - c0
        elif self._graph is None:
            # one eager step configures kernel attributes and warms the allocator, then capture
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

    def train_step(self, idx: torch.Tensor, targets: torch.Tensor, attack_scale: float = 1.0, **_ignored) -> float:
        """Public step: ``idx``/``targets`` may live on the host (pinned or not) or the device.
        Returns the mean loss (one 4-byte D2H read).  ``attack_scale`` != 1 makes this rank poison its
        gradient bucket (scale / sign flip) before aggregation — the fault-injection path of the attacker."""
        if attack_scale != self._attack_scale_host:
            self.attack_scale.fill_(float(attack_scale))
            self._attack_scale_host = float(attack_scale)
        if idx.device.type == "cpu":
            self.idx_host.copy_(idx.view(self.B, self.T))
            self.tgt_host.copy_(targets.view(self.B, self.T))
            self.idx[0].copy_(self.idx_host, non_blocking=True)
            self.tgt[0].copy_(self.tgt_host, non_blocking=True)
        else:
            self.idx[0].copy_(idx.view(self.B, self.T))
            self.tgt[0].copy_(targets.view(self.B, self.T))
        self.run_step_device()
        self.loss_host.copy_(self.loss_sum, non_blocking=True)
        if self.security:
            self.act_stats_host.copy_(self.act_stats, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(self.loss_host.item()) / self.M

    def kernels_per_step(self) -> int:
        """Kernel-library calls issued by one step, counted while the step was last built (eagerly or under graph
        capture); every call launches at least one kernel of ours (the statistics call launches seven)."""
        if self._calls_per_step is None:
            raise RuntimeError("run at least two steps first (the count is taken when the step is built)")
        return int(self._calls_per_step)

    def set_security(self, enabled: bool) -> None:
        """Toggle the in-step security kernels (drops the captured graph; used for overhead runs).  With several
        ranks the aggregation switches between K2 (trust-weighted, statistics fused) and K4 (plain mean all-reduce)."""
        if enabled != self.security:
            self.security = enabled
            self._graph = None

    def security_snapshot(self) -> Dict[str, torch.Tensor]:
        """Host copies of the device-resident security state (one small sync)."""
        r = self.reducer
        out = {"trust": r.trust.cpu(), "status": r.status.cpu(), "stats": r.stats.cpu(), "coeff": r.coeff.cpu(),
               "flags": r.flags.cpu(), "act_stats": self.act_stats.cpu()}
        if self.zero:
            out["param_flags"] = self.pg_flags.cpu()
            out["param_drift"] = self.pg_drift.sum(0).sqrt().cpu()
        return out

    # ------------------------------------------------------------------ inference-style forward (tests, validation)
    @torch.no_grad()
    def forward_logits(self, idx: torch.Tensor) -> torch.Tensor:
        assert self.has_embed and self.has_head
        self.idx[0].copy_(idx.view(self.B, self.T))
        K.embedding_fwd(self.idx[0], self.w16("wte"), self.w16("wpe"), self.res[0, 0], self.T)
        self._forward_blocks(0)
        K.linear_fwd(self.lnf[0], self.w16(self.head_name), self.logits)
        return self.logits.view(self.B, self.T, self.Vp)[..., : self.V].float()

    @torch.no_grad()
    def head_loss_sum(self, slot: int, targets: torch.Tensor) -> torch.Tensor:
        """Summed cross-entropy (device double) of the tokens whose final LayerNorm output ``_forward(slot)`` just produced;
        forward-only (validation of the stage that holds the head)."""
        assert self.has_head
        K.linear_fwd(self.lnf[slot], self.w16(self.head_name), self.logits)
        lg = self.logits[:, : self.V].float()
        return torch.nn.functional.cross_entropy(lg, targets.reshape(-1), reduction="sum").double()

    @torch.no_grad()
    def eval_loss(self, idx: torch.Tensor, targets: torch.Tensor) -> float:
        """Mean cross-entropy of one batch with the engine's *current* parameters (no gradient, no update)."""
        logits = self.forward_logits(idx)
        return float(torch.nn.functional.cross_entropy(logits.view(-1, self.V), targets.view(-1).to(self.device)).item())
