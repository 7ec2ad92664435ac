"""Fused GPT-2 training engine for B200 (sm_100a).

This is the hot path behind ``DistributedTrainer`` when the model is GPT-2 on a GPU and behind
``bench.py``.  It replaces what the reference would run through torch eager / cuBLAS / ATen
(SURVEY 2.7, rows "stage forward", "loss", "backward", "optimizer", "gradient aggregation",
"per-stage stats") with an explicit forward/backward schedule over hand-written kernels:

* every GEMM (QKV, attention-out, MLP up/down, LM head; their dgrads and wgrads) is the tcgen05
  kernel in csrc/gemm_tcgen05.cu — bias, GELU (+saved pre-activation), residual add and the GELU
  derivative are epilogues; dgrad/wgrad consume the stored operands MN-major (no transposes);
  wgrads accumulate split-K into an fp32 main-grad buffer,
* attention is csrc/attention.cu (tcgen05 S=QK^T / PV / dK,dV / dQ, TMEM accumulators),
* LayerNorm fwd/bwd (bwd fused with the residual-gradient add), bias grads, embedding, fused
  cross-entropy and fused AdamW are csrc/elementwise.cu,
* security runs inside the step on the device: activation statistics (csrc/stats.cu), bucket
  pack + norm, trust-weighted reduction with fused median/cosine statistics over peer memory,
  device-side coefficient + trust state machine (csrc/trust_allreduce.cu),
* the whole step is captured once into a CUDA graph and replayed; parameters, optimizer state,
  gradients and activations live in preallocated flat buffers sized for HBM3e (no allocator
  traffic on the step path).

Parameter layout: one flat fp32 master buffer (+ bf16 working copy, Adam moments, fp32 main
grads, bf16 gradient bucket).  Weight-decayed tensors come first so AdamW needs no per-tensor
table.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch

from ..utils.timing import nvtx_range
from ..models.gpt2 import GPT2Config
from ..ops import _lib, nn as K
from ..ops.stats import tensor_statistics_device
from ..parallel.symm import SymmetricTrustReducer

__all__ = ["GPT2Engine"]


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class GPT2Engine:
    def __init__(self, cfg: GPT2Config, batch_size: int, seq_len: int, lr: float = 5e-5, weight_decay: float = 0.01,
                 betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8, device: Optional[torch.device] = None,
                 security: bool = True, use_graph: bool = True, seed: int = 0, clip_ratio: float = 2.0,
                 trust_threshold: float = 0.7):
        assert cfg.n_embd % cfg.n_head == 0 and cfg.n_embd // cfg.n_head == 64, "engine is specialised for head_dim 64"
        assert seq_len % 128 == 0 and seq_len <= cfg.n_positions
        self.cfg = cfg
        self.B, self.T = batch_size, seq_len
        self.M = batch_size * seq_len
        self.C, self.L, self.H = cfg.n_embd, cfg.n_layer, cfg.n_head
        self.V, self.Vp = cfg.vocab_size, cfg.padded_vocab_size
        self.lr, self.wd, self.betas, self.eps = lr, weight_decay, betas, eps
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.security = security
        self.use_graph = use_graph
        self.step_count = 0
        self._graph = None
        self._build_params(seed, trust_threshold, clip_ratio)
        self._build_activations()

    # ------------------------------------------------------------------ parameters
    def _param_specs(self) -> List[Tuple[str, Tuple[int, ...], bool]]:
        C, L = self.C, self.L
        specs: List[Tuple[str, Tuple[int, ...], bool]] = [("wte", (self.Vp, C), True), ("wpe", (self.cfg.n_positions, C), True)]
        for l in range(L):
            specs += [(f"h{l}.qkv_w", (3 * C, C), True), (f"h{l}.o_w", (C, C), True),
                      (f"h{l}.fc_w", (4 * C, C), True), (f"h{l}.proj_w", (C, 4 * C), True)]
        for l in range(L):
            specs += [(f"h{l}.ln1_w", (C,), False), (f"h{l}.ln1_b", (C,), False), (f"h{l}.qkv_b", (3 * C,), False),
                      (f"h{l}.o_b", (C,), False), (f"h{l}.ln2_w", (C,), False), (f"h{l}.ln2_b", (C,), False),
                      (f"h{l}.fc_b", (4 * C,), False), (f"h{l}.proj_b", (C,), False)]
        specs += [("lnf_w", (C,), False), ("lnf_b", (C,), False)]
        return specs

    def _build_params(self, seed: int, trust_threshold: float, clip_ratio: float) -> None:
        dev = self.device
        specs = self._param_specs()
        self.offsets: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        off = 0
        self.n_decay = 0
        for name, shape, decay in specs:
            n = math.prod(shape)
            assert n % 8 == 0
            self.offsets[name] = (off, shape)
            off += n
            if decay:
                self.n_decay = off
        self.P = _round_up(off, 64)
        self.p32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.p16 = torch.zeros(self.P, dtype=torch.bfloat16, device=dev)
        self.m = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.v = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.g32 = torch.zeros(self.P, dtype=torch.float32, device=dev)
        self.reducer = SymmetricTrustReducer(self.P, torch.bfloat16, dev, clip_ratio=clip_ratio,
                                             trust_threshold=trust_threshold)
        self.bucket = self.reducer.bucket.tensor[: self.P]
        self.world, self.rank = self.reducer.world, self.reducer.rank
        self.hyper = torch.zeros(4, dtype=torch.float32, device=dev)
        self.attack_scale = torch.ones(1, dtype=torch.float32, device=dev)      # fault-injection hook (1 = honest)
        self._attack_scale_host = 1.0
        # GPT-2 initialisation (same on every rank: replicas start identical)
        g = torch.Generator(device="cpu").manual_seed(seed)
        for name, (o, shape) in self.offsets.items():
            n = math.prod(shape)
            view = self.p32[o:o + n]
            if name.endswith("_w") and len(shape) == 1:
                view.fill_(1.0)
            elif len(shape) == 1:
                view.zero_()
            else:
                std = 0.02 / math.sqrt(2 * self.L) if name.endswith(("o_w", "proj_w")) else 0.02
                view.copy_((torch.randn(n, generator=g) * std).to(dev))
        if self.Vp > self.V:
            o, shape = self.offsets["wte"]
            self.p32[o + self.V * self.C:o + self.Vp * self.C].zero_()
        K.cast_f32_bf16(self.p32, self.p16)

    def w32(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.p32[o:o + math.prod(shape)].view(shape)

    def w16(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.p16[o:o + math.prod(shape)].view(shape)

    def grad32(self, name: str) -> torch.Tensor:
        o, shape = self.offsets[name]
        return self.g32[o:o + math.prod(shape)].view(shape)

    # ------------------------------------------------------------------ module interop
    _MODULE_MAP = {"qkv_w": "attn.c_attn.weight", "qkv_b": "attn.c_attn.bias", "o_w": "attn.c_proj.weight",
                   "o_b": "attn.c_proj.bias", "fc_w": "mlp.c_fc.weight", "fc_b": "mlp.c_fc.bias",
                   "proj_w": "mlp.c_proj.weight", "proj_b": "mlp.c_proj.bias", "ln1_w": "ln_1.weight",
                   "ln1_b": "ln_1.bias", "ln2_w": "ln_2.weight", "ln2_b": "ln_2.bias"}

    def _module_tensors(self, module) -> Dict[str, torch.Tensor]:
        out = {"wte": module.transformer.wte.weight, "wpe": module.transformer.wpe.weight,
               "lnf_w": module.transformer.ln_f.weight, "lnf_b": module.transformer.ln_f.bias}
        for l, blk in enumerate(module.transformer.h):
            sd = dict(blk.named_parameters())
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
        K.cast_f32_bf16(self.p32, self.p16)

    @torch.no_grad()
    def export_to_module(self, module) -> None:
        for name, t in self._module_tensors(module).items():
            t.copy_(self.w32(name).to(t.device, t.dtype))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return {"p32": self.p32.clone(), "m": self.m.clone(), "v": self.v.clone(), "hyper": self.hyper.clone(),
                "step": torch.tensor(self.step_count), "trust": self.reducer.trust.clone(),
                "status": self.reducer.status.clone()}

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        self.p32.copy_(sd["p32"]); self.m.copy_(sd["m"]); self.v.copy_(sd["v"]); self.hyper.copy_(sd["hyper"])
        self.step_count = int(sd["step"])
        self.reducer.trust.copy_(sd["trust"]); self.reducer.status.copy_(sd["status"])
        K.cast_f32_bf16(self.p32, self.p16)

    # ------------------------------------------------------------------ activations
    def _build_activations(self) -> None:
        dev, M, C, L = self.device, self.M, self.C, self.L
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.idx = torch.zeros(self.B, self.T, dtype=torch.int64, device=dev)
        self.tgt = torch.zeros(self.B, self.T, dtype=torch.int64, device=dev)
        self.res = torch.empty(L + 1, M, C, **bf)
        self.ln1 = torch.empty(L, M, C, **bf)
        self.qkv = torch.empty(L, M, 3 * C, **bf)
        self.atty = torch.empty(L, M, C, **bf)
        self.x2 = torch.empty(L, M, C, **bf)
        self.ln2 = torch.empty(L, M, C, **bf)
        self.fch = torch.empty(L, M, 4 * C, **bf)       # gelu'(fc pre-activation), consumed by the MLP backward
        self.fcg = torch.empty(L, M, 4 * C, **bf)
        self.ln_mean = torch.empty(2 * L + 1, M, **f32)
        self.ln_rstd = torch.empty(2 * L + 1, M, **f32)
        self.lse = torch.empty(L, self.B * self.H * self.T, **f32)
        self.lnf = torch.empty(M, C, **bf)
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
        self.act_stats = torch.zeros(12, dtype=torch.float64, device=dev)
        self.act_stats_host = torch.zeros(12, dtype=torch.float64).pin_memory()
        self.sec_host = torch.zeros(self.world * 6, dtype=torch.float32).pin_memory()
        self.idx_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.tgt_host = torch.zeros(self.B, self.T, dtype=torch.int64).pin_memory()
        self.loss_host = torch.zeros(1, dtype=torch.float32).pin_memory()

    def _split_k(self, n_out: int, k_in: int) -> int:
        """0 = let the kernel library choose the split that fills whole waves of the persistent grid."""
        return 0

    # ------------------------------------------------------------------ forward / backward
    def _forward(self) -> None:
        K.embedding_fwd(self.idx, self.w16("wte"), self.w16("wpe"), self.res[0], self.T)
        self._forward_layers_only()
        K.xent_fwd_bwd(self.logits, self.tgt.view(-1), self.row_loss, self.loss_sum, self.V, 1.0 / self.M)

    def _backward(self) -> None:
        L, M, C = self.L, self.M, self.C
        dlogits = self.logits
        # LM head (tied): dlnf = dlogits @ wte ; dwte += dlogits^T @ lnf
        K.linear_dgrad(dlogits, self.w16("wte"), self.dln)
        K.linear_wgrad(dlogits, self.lnf, self.grad32("wte"), split_k=1)
        dx, dx_next = self.dx_a, self.dx_b
        # every LayerNorm backward also emits the column sums of the dx it writes = the bias gradient of the linear layer
        # below it (MLP down-projection under ln_f / ln_1, attention output projection under ln_2)
        K.layernorm_bwd(self.dln, self.res[L], self.w32("lnf_w"), self.ln_mean[2 * L], self.ln_rstd[2 * L], None, dx,
                        self.grad32("lnf_w"), self.grad32("lnf_b"), dxsum=self.grad32(f"h{L - 1}.proj_b"))
        for l in reversed(range(L)):
            p = f"h{l}."
            # ---- MLP ----
            K.linear_wgrad(dx, self.fcg[l], self.grad32(p + "proj_w"), split_k=self._split_k(C, 4 * C))
            K.linear_dgrad(dx, self.w16(p + "proj_w"), self.dfch, epilogue=K.EPI_MUL, aux_in=self.fch[l],
                           colsum=self.grad32(p + "fc_b"))          # fc bias grad from the epilogue panels
            K.linear_wgrad(self.dfch, self.ln2[l], self.grad32(p + "fc_w"), split_k=self._split_k(4 * C, C))
            K.linear_dgrad(self.dfch, self.w16(p + "fc_w"), self.dln)
            K.layernorm_bwd(self.dln, self.x2[l], self.w32(p + "ln2_w"), self.ln_mean[2 * l + 1], self.ln_rstd[2 * l + 1],
                            dx, dx_next, self.grad32(p + "ln2_w"), self.grad32(p + "ln2_b"), dxsum=self.grad32(p + "o_b"))
            dx, dx_next = dx_next, dx
            # ---- attention ----
            K.linear_wgrad(dx, self.atty[l], self.grad32(p + "o_w"), split_k=self._split_k(C, C))
            K.linear_dgrad(dx, self.w16(p + "o_w"), self.datty)
            K.attention_bwd(self.qkv[l], self.atty[l], self.datty, self.lse[l], self.delta, self.dqkv, self.B, self.T, self.H,
                            dbias=self.grad32(p + "qkv_b"))         # qkv bias grad from the accumulator write-out
            K.linear_wgrad(self.dqkv, self.ln1[l], self.grad32(p + "qkv_w"), split_k=self._split_k(3 * C, C))
            K.linear_dgrad(self.dqkv, self.w16(p + "qkv_w"), self.dln)
            K.layernorm_bwd(self.dln, self.res[l], self.w32(p + "ln1_w"), self.ln_mean[2 * l], self.ln_rstd[2 * l],
                            dx, dx_next, self.grad32(p + "ln1_w"), self.grad32(p + "ln1_b"),
                            dxsum=self.grad32(f"h{l - 1}.proj_b") if l > 0 else None)
            dx, dx_next = dx_next, dx
        K.embedding_bwd(self.idx, dx, self.grad32("wte"), self.grad32("wpe"), self.T)

    def _security_and_update(self) -> None:
        r = self.reducer
        if self.security:
            tensor_statistics_device(self.res[self.L].view(-1), out=self.act_stats)       # K5 on the last residual
        K.pack_norm(self.g32, self.bucket, r.sumsq, self.attack_scale)                    # K1 (+ fault-injection scale)
        if self.security or self.world > 1:
            r.enqueue_reduce(update_trust=self.security)                                  # K2 + K3
        K.adam_tick(self.hyper, self.lr, self.betas[0], self.betas[1])
        K.adamw(self.p32, self.p16, self.bucket, self.m, self.v, self.n_decay, self.lr, self.betas[0], self.betas[1],
                self.eps, self.wd, 1, hyper=self.hyper)                                   # K14

    def _step_body(self) -> None:
        self.loss_sum.zero_()
        self.g32.zero_()
        with nvtx_range("engine.forward"):
            self._forward()
        with nvtx_range("engine.backward"):
            self._backward()
        with nvtx_range("engine.security+allreduce+adamw"):
            self._security_and_update()

    # ------------------------------------------------------------------ public step
    def run_step_device(self) -> None:
        """Enqueue one full training step on whatever is in ``self.idx`` / ``self.tgt`` (no host sync)."""
        if not self.use_graph:
            self._step_body()
        elif self._graph is None:
            # two eager steps configure kernel attributes and warm the allocator, then capture
            if self.step_count == 0:
                self._step_body()
                torch.cuda.synchronize()
                self.step_count += 1
                return
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    self._step_body()
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
            self.idx.copy_(self.idx_host, non_blocking=True)
            self.tgt.copy_(self.tgt_host, non_blocking=True)
        else:
            self.idx.copy_(idx.view(self.B, self.T))
            self.tgt.copy_(targets.view(self.B, self.T))
        self.run_step_device()
        self.loss_host.copy_(self.loss_sum, non_blocking=True)
        if self.security:
            self.act_stats_host.copy_(self.act_stats, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(self.loss_host.item()) / self.M

    def kernels_per_step(self) -> int:
        """Number of kernel launches of ours in one step (library calls x kernels per call)."""
        L = self.L
        P = 2                                     # radix-select passes for bf16
        n = 1 + L * 7 + 2 + 1                     # embedding, per-layer fwd (2 LN, 4 GEMM, attn), lnf+head, xent
        n += 2 + 1 + L * (4 + 4 + 2 + 2) + 1      # head dgrad/wgrad, lnf bwd, per layer: 4 wgrad, 4 dgrad, 2 LN, 2 attn (bias grads and delta fused), emb
        n += 1 + 2                                # pack+norm, adam tick, adamw
        if self.security:
            n += 2 + 2 * P + 1 + 5                # stats (moments, final, P x (hist, select), finalize) + exchange/coeff/reduce/exchange/finalize
        elif self.world > 1:
            n += 5
        return n

    def set_security(self, enabled: bool) -> None:
        """Toggle the in-step security kernels (drops the captured graph; used for overhead runs)."""
        if enabled != self.security:
            self.security = enabled
            self._graph = None

    def security_snapshot(self) -> Dict[str, torch.Tensor]:
        """Host copies of the device-resident security state (one small sync)."""
        r = self.reducer
        return {"trust": r.trust.cpu(), "status": r.status.cpu(), "stats": r.stats.cpu(), "coeff": r.coeff.cpu(),
                "flags": r.flags.cpu(), "act_stats": self.act_stats.cpu()}

    # ------------------------------------------------------------------ inference-style forward (tests)
    @torch.no_grad()
    def forward_logits(self, idx: torch.Tensor) -> torch.Tensor:
        self.idx.copy_(idx.view(self.B, self.T))
        self.tgt.zero_()
        K.embedding_fwd(self.idx, self.w16("wte"), self.w16("wpe"), self.res[0], self.T)
        self._forward_layers_only()
        return self.logits.view(self.B, self.T, self.Vp)[..., : self.V].float()

    def _forward_layers_only(self) -> None:
        L = self.L
        for l in range(L):
            p = f"h{l}."
            K.layernorm_fwd(self.res[l], self.w32(p + "ln1_w"), self.w32(p + "ln1_b"), self.ln1[l],
                            self.ln_mean[2 * l], self.ln_rstd[2 * l], self.cfg.layer_norm_eps)
            K.linear_fwd(self.ln1[l], self.w16(p + "qkv_w"), self.qkv[l], bias=self.w32(p + "qkv_b"))
            K.attention_fwd(self.qkv[l], self.atty[l], self.lse[l], self.B, self.T, self.H)
            K.linear_fwd(self.atty[l], self.w16(p + "o_w"), self.x2[l], bias=self.w32(p + "o_b"),
                         epilogue=K.EPI_RESIDUAL, aux_in=self.res[l])
            K.layernorm_fwd(self.x2[l], self.w32(p + "ln2_w"), self.w32(p + "ln2_b"), self.ln2[l],
                            self.ln_mean[2 * l + 1], self.ln_rstd[2 * l + 1], self.cfg.layer_norm_eps)
            K.linear_fwd(self.ln2[l], self.w16(p + "fc_w"), self.fcg[l], bias=self.w32(p + "fc_b"),
                         epilogue=K.EPI_GELU_DG, aux_out=self.fch[l])     # fch holds gelu'(pre): backward multiplies
            K.linear_fwd(self.fcg[l], self.w16(p + "proj_w"), self.res[l + 1], bias=self.w32(p + "proj_b"),
                         epilogue=K.EPI_RESIDUAL, aux_in=self.x2[l])
        K.layernorm_fwd(self.res[L], self.w32("lnf_w"), self.w32("lnf_b"), self.lnf, self.ln_mean[2 * L],
                        self.ln_rstd[2 * L], self.cfg.layer_norm_eps)
        K.linear_fwd(self.lnf, self.w16("wte"), self.logits)
