"""GPT-2 (small / medium / large / tiny-test) as a chain of tensor->tensor units.

The reference expects ``ModelFactory().create_model('gpt2')`` to return something with
``.transformer.h`` (distributed_trainer.py:126,133) and chains the blocks with ``nn.Sequential``
— which cannot work with HF ``GPT2Block`` (tuple outputs) and leaves embeddings / ``ln_f`` /
head on no stage.  Here every unit maps one tensor to one tensor, so any
contiguous slice of ``stage_units()`` is a valid pipeline stage: the first unit owns the
embeddings, the last owns ``ln_f`` and the (tied) LM head.

This nn.Module is the portable definition (CPU, gloo tests, parity oracle).  On a B200 the
same parameters are executed by :class:`trustworthy_dl.engine.gpt2_engine.GPT2Engine`, which
runs hand-written sm_100a kernels (tcgen05 GEMMs with fused epilogues, flash attention,
LayerNorm, fused CE and AdamW) over one flat parameter buffer.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List

import torch
import torch.nn as nn
import torch.nn.functional as F

__all__ = ["GPT2Config", "GPT2LMHeadModel", "GPT2Block", "GPT2Embed", "GPT2Head", "GPT2_PRESETS"]


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    padded_vocab_size: int = 50304      # multiple of 128 so the LM-head GEMM tiles evenly
    layer_norm_eps: float = 1e-5

    def __post_init__(self):
        if self.padded_vocab_size < self.vocab_size:
            self.padded_vocab_size = (self.vocab_size + 127) // 128 * 128


GPT2_PRESETS = {
    "gpt2": dict(n_embd=768, n_layer=12, n_head=12),
    "gpt2-small": dict(n_embd=768, n_layer=12, n_head=12),
    "gpt2-medium": dict(n_embd=1024, n_layer=24, n_head=16),
    "gpt2-large": dict(n_embd=1280, n_layer=36, n_head=20),
    "gpt2-tiny": dict(n_embd=128, n_layer=4, n_head=2, n_positions=128, vocab_size=512, padded_vocab_size=512),
}


class CausalSelfAttention(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.n_head = cfg.n_head
        self.c_attn = nn.Linear(cfg.n_embd, 3 * cfg.n_embd)
        self.c_proj = nn.Linear(cfg.n_embd, cfg.n_embd)

    def forward(self, x):
        B, T, C = x.shape
        q, k, v = self.c_attn(x).split(C, dim=2)
        q, k, v = (t.view(B, T, self.n_head, C // self.n_head).transpose(1, 2) for t in (q, k, v))
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.c_proj(y.transpose(1, 2).reshape(B, T, C))


class MLP(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.c_fc = nn.Linear(cfg.n_embd, 4 * cfg.n_embd)
        self.c_proj = nn.Linear(4 * cfg.n_embd, cfg.n_embd)

    def forward(self, x):
        return self.c_proj(F.gelu(self.c_fc(x), approximate="tanh"))


class GPT2Block(nn.Module):
    """Pre-LN transformer block; tensor in, tensor out."""

    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.ln_1 = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
        self.attn = CausalSelfAttention(cfg)
        self.ln_2 = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
        self.mlp = MLP(cfg)

    def forward(self, x):
        x = x + self.attn(self.ln_1(x))
        return x + self.mlp(self.ln_2(x))


class GPT2Embed(nn.Module):
    def __init__(self, wte: nn.Embedding, wpe: nn.Embedding):
        super().__init__()
        self.wte, self.wpe = wte, wpe

    def forward(self, idx):
        T = idx.shape[1]
        pos = torch.arange(T, device=idx.device)
        return self.wte(idx) + self.wpe(pos)[None]


class GPT2Head(nn.Module):
    """ln_f + LM head over the padded vocabulary (logits of the pad columns are sliced off)."""

    def __init__(self, ln_f: nn.LayerNorm, wte: nn.Embedding, vocab_size: int, tie: bool = True):
        super().__init__()
        self.ln_f = ln_f
        self.vocab_size = vocab_size
        if tie:
            self._tied = [wte]                       # not registered twice
            self.weight = None
        else:
            self._tied = None
            self.weight = nn.Parameter(wte.weight.detach().clone())

    def forward(self, x):
        w = self._tied[0].weight if self._tied is not None else self.weight
        return F.linear(self.ln_f(x), w)[..., : self.vocab_size]


class _Transformer(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.wte = nn.Embedding(cfg.padded_vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList(GPT2Block(cfg) for _ in range(cfg.n_layer))
        self.ln_f = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)


class GPT2LMHeadModel(nn.Module):
    task = "lm"

    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.config = cfg
        self.transformer = _Transformer(cfg)
        self.embed = GPT2Embed(self.transformer.wte, self.transformer.wpe)
        self.head = GPT2Head(self.transformer.ln_f, self.transformer.wte, cfg.vocab_size, tie=True)
        self.apply(self._init)
        for name, p in self.named_parameters():
            if name.endswith("c_proj.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * cfg.n_layer))

    @staticmethod
    def _init(m):
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def forward(self, idx):
        x = self.embed(idx)
        for blk in self.transformer.h:
            x = blk(x)
        return self.head(x)

    def stage_units(self) -> List[nn.Module]:
        """Ordered tensor->tensor units; any contiguous grouping is a valid stage split."""
        return [self.embed, *self.transformer.h, self.head]

    def untie_head(self) -> None:
        """Give the head its own weight (needed when embed and head live on different ranks)."""
        if self.head._tied is not None:
            self.head.weight = nn.Parameter(self.transformer.wte.weight.detach().clone())
            self.head._tied = None

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())

    def loss(self, logits, targets):
        return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), targets.reshape(-1))
