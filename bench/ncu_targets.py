#!/usr/bin/env python
"""One pass over every hot kernel at the shapes of the GPT-2-medium step (B = 32, T = 1024), for `ncu -k regex:<kernel>`
captures (scripts/ncu_capture.sh).  No arguments, so the plain run and the profiled runs share one command line.
The cross-GPU reduction is run with 8 ranks emulated on this GPU (local pointers): instruction mix and memory behaviour
of the kernel, not NVLink counters (ncu never wraps a multi-rank command here)."""
from __future__ import annotations

import ctypes
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch

from trustworthy_dl.ops import _lib, nn as K
from trustworthy_dl.ops._lib import i32, i64, ptr, stream_ptr
from trustworthy_dl.ops.stats import tensor_statistics_device


def main() -> int:
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    B, T, C, H, V, Vp = 32, 1024, 1024, 16, 50257, 50304
    M = B * T
    bf = torch.bfloat16
    reps = 2
    x = torch.randn(M, C, device=dev, dtype=bf)
    # ---- GEMMs: plain (f_qkv), two outputs (f_fc: GELU + GELU'), second input + column sums (d_proj), split-K wgrad (w_fc)
    w_qkv = torch.randn(3 * C, C, device=dev, dtype=bf) * 0.02
    b_qkv = torch.zeros(3 * C, device=dev)
    qkv = torch.empty(M, 3 * C, device=dev, dtype=bf)
    w_fc = torch.randn(4 * C, C, device=dev, dtype=bf) * 0.02
    b_fc = torch.zeros(4 * C, device=dev)
    fcg = torch.empty(M, 4 * C, device=dev, dtype=bf); fch = torch.empty(M, 4 * C, device=dev, dtype=bf)
    w_proj = torch.randn(C, 4 * C, device=dev, dtype=bf) * 0.02
    dfch = torch.empty(M, 4 * C, device=dev, dtype=bf)
    cs = torch.zeros(4 * C, device=dev)
    dw = torch.zeros(4 * C, C, device=dev)
    for _ in range(reps):
        K.linear_fwd(x, w_qkv, qkv, bias=b_qkv)
        K.linear_fwd(x, w_fc, fcg, bias=b_fc, epilogue=K.EPI_GELU_DG, aux_out=fch)
        K.linear_dgrad(x, w_proj, dfch, epilogue=K.EPI_MUL, aux_in=fch, colsum=cs)
        K.linear_wgrad(dfch, x, dw, split_k=0)
    # ---- attention
    o = torch.empty(M, C, device=dev, dtype=bf); lse = torch.empty(B * H * T, device=dev)
    do = torch.randn_like(o); dqkv = torch.empty_like(qkv); delta = torch.empty(B * H * T, device=dev)
    dbias = torch.zeros(3 * C, device=dev)
    qkv.normal_()
    for _ in range(reps):
        K.attention_fwd(qkv, o, lse, B, T, H)
        K.attention_bwd(qkv, o, do, lse, delta, dqkv, B, T, H, dbias=dbias)
    # ---- LayerNorm, cross-entropy, AdamW, statistics
    w = torch.ones(C, device=dev); b = torch.zeros(C, device=dev)
    mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
    y = torch.empty_like(x); dx = torch.empty_like(x)
    gw = torch.zeros(C, device=dev); gb = torch.zeros(C, device=dev); dxs = torch.zeros(C, device=dev)
    logits = torch.randn(M, Vp, device=dev, dtype=bf)
    tgt = torch.randint(0, V, (M,), device=dev)
    row_loss = torch.empty(M, device=dev); loss_sum = torch.zeros(1, device=dev)
    P = 75_000_000 // 1024 * 1024                    # one gradient segment of GPT-2 medium (6 blocks)
    p32 = torch.randn(P, device=dev) * 0.02; p16 = torch.empty(P, device=dev, dtype=bf)
    g32 = torch.randn(P, device=dev) * 1e-3; g16 = torch.empty(P, device=dev, dtype=bf)
    m1 = torch.zeros(P, device=dev); v1 = torch.zeros(P, device=dev)
    sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
    stats = torch.zeros(12, dtype=torch.float64, device=dev)
    for _ in range(reps):
        K.layernorm_fwd(x, w, b, y, mean, rstd)
        K.layernorm_bwd(y, x, w, mean, rstd, x, dx, gw, gb, dxsum=dxs)
        K.xent_fwd_bwd(logits, tgt, row_loss, loss_sum, V, 1.0 / M)
        K.pack_norm(g32, g16, sumsq, None)
        K.adamw(p32, p16, g16, m1, v1, P // 2, 5e-5, 0.9, 0.999, 1e-8, 0.01, 3)
        tensor_statistics_device(x.view(-1), out=stats)
    # ---- trust-weighted reduction, 8 emulated ranks (each "rank" = one launch over its slice; local pointers)
    W, N = 8, 8 * 8 * 1024 * 1024
    bufs = [(torch.randn(N, device=dev) * 1e-3).to(bf) for _ in range(W)]
    outs = [torch.empty(N, device=dev, dtype=bf) for _ in range(W)]
    arr_in = (ctypes.c_void_p * W)(*[t.data_ptr() for t in bufs])
    arr_out = (ctypes.c_void_p * W)(*[t.data_ptr() for t in outs])
    coeff = torch.full((W,), 1.0 / W, device=dev)
    partial = torch.zeros(W + 1, dtype=torch.float64, device=dev)
    nvec = N // 8
    per = nvec // W
    for r in range(W):
        _lib.call("tdl_trust_reduce", arr_in, arr_out, i32(W), i32(1), i64(r * per), i64((r + 1) * per), ptr(coeff), ptr(partial),
                  stream_ptr())
    torch.cuda.synchronize()
    print("ok")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
