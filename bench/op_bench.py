#!/usr/bin/env python
"""Device-timed micro-benchmarks of the non-GEMM hot ops at GPT-2-medium step shapes (B=32, T=1024), L2 flushed
between iterations.  One line per op: time, bytes moved (algorithmic), achieved GB/s and fraction of the measured HBM
copy bandwidth in MEASURED_PEAKS.json.  The flush leaves up to one L2 (126 MB) of dirty lines that are written back
while the timed kernel runs, so kernels that move only ~100-300 MB (LayerNorm, bias gradients) read low here; use the
numbers to compare variants of one op, and the `ncu` launch lists for absolute times of the small kernels.

  python bench/op_bench.py [--ops xent,ln_bwd,ln_fwd,bias_grad,dproj,attn] [--iters 10]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch

from trustworthy_dl.ops import nn as K


def timed(fn, iters, flush):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.add_(1.0)                                    # 512 MB write: evicts the 126 MB L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ops", default="xent,ln_fwd,ln_bwd,bias_grad,dproj,attn")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--seq", type=int, default=1024)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6548.0))
    M, C, T, H, V, Vp = args.batch * args.seq, 1024, args.seq, 16, 50257, 50304
    flush = torch.zeros(128 * 1024 * 1024, device=dev)
    bf = torch.bfloat16
    rows = []

    def report(name, us, nbytes, note=""):
        gbps = nbytes / us / 1e3
        rows.append((name, us, nbytes / 1e6, gbps, gbps / hbm))
        print(f"{name:28s} {us:9.1f} us  {nbytes / 1e6:9.1f} MB  {gbps:8.1f} GB/s  {gbps / hbm:5.2f} of HBM copy  {note}", flush=True)

    ops = args.ops.split(",")
    if "xent" in ops:
        logits = torch.randn(M, Vp, device=dev, dtype=bf)
        keep = logits.clone()
        tgt = torch.randint(0, V, (M,), device=dev)
        row_loss = torch.empty(M, device=dev)
        loss_sum = torch.zeros(1, device=dev)

        def f():
            K.xent_fwd_bwd(logits, tgt, row_loss, loss_sum, V, 1.0 / M)
        us = timed(f, args.iters, flush)
        report("xent fwd+bwd (in place)", us, 2 * M * Vp * 2, "1 read + 1 write of the logits")
        del logits, keep
    x = torch.randn(M, C, device=dev, dtype=bf)
    w = torch.ones(C, device=dev); b = torch.zeros(C, device=dev)
    mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
    y = torch.empty_like(x)
    if "ln_fwd" in ops:
        us = timed(lambda: K.layernorm_fwd(x, w, b, y, mean, rstd), args.iters, flush)
        report("layernorm fwd", us, 2 * M * C * 2)
    if "ln_bwd" in ops:
        K.layernorm_fwd(x, w, b, y, mean, rstd)
        dy = torch.randn_like(x); dres = torch.randn_like(x); dx = torch.empty_like(x)
        dw = torch.zeros(C, device=dev); db = torch.zeros(C, device=dev)
        us = timed(lambda: K.layernorm_bwd(dy, x, w, mean, rstd, dres, dx, dw, db), args.iters, flush)
        report("layernorm bwd (+residual)", us, 4 * M * C * 2)
    if "bias_grad" in ops:
        for N in (C, 3 * C, 4 * C):
            g = torch.randn(M, N, device=dev, dtype=bf)
            db = torch.zeros(N, device=dev)
            us = timed(lambda: K.bias_grad(g, db), args.iters, flush)
            report(f"bias_grad N={N}", us, M * N * 2)
            del g
    if "dproj" in ops:
        dyp = torch.randn(M, C, device=dev, dtype=bf)
        wp = torch.randn(C, 4 * C, device=dev, dtype=bf) * 0.02
        pre = torch.randn(M, 4 * C, device=dev, dtype=bf)
        out = torch.empty(M, 4 * C, device=dev, dtype=bf)
        cs = torch.zeros(4 * C, device=dev)
        fl = 2.0 * M * 4 * C * C
        us0 = timed(lambda: K.linear_dgrad(dyp, wp, out, epilogue=K.EPI_DGELU, aux_in=pre), args.iters, flush)
        report("d_proj GEMM (dGELU)", us0, (M * C + 2 * M * 4 * C) * 2, f"{fl / us0 / 1e6:.0f} TFLOP/s")
        us1 = timed(lambda: K.linear_dgrad(dyp, wp, out, epilogue=K.EPI_DGELU, aux_in=pre, colsum=cs), args.iters, flush)
        report("d_proj GEMM (dGELU+colsum)", us1, (M * C + 2 * M * 4 * C) * 2, f"{fl / us1 / 1e6:.0f} TFLOP/s; fused bias grad costs {us1 - us0:+.1f} us")
        usm = timed(lambda: K.linear_dgrad(dyp, wp, out, epilogue=K.EPI_MUL, aux_in=pre, colsum=cs), args.iters, flush)
        report("d_proj GEMM (*saved gelu'+colsum)", usm, (M * C + 2 * M * 4 * C) * 2, f"{fl / usm / 1e6:.0f} TFLOP/s")
        del dyp, wp, pre, out
        xin = torch.randn(M, C, device=dev, dtype=bf)
        wfc = torch.randn(4 * C, C, device=dev, dtype=bf) * 0.02
        bfc = torch.zeros(4 * C, device=dev)
        pre = torch.empty(M, 4 * C, device=dev, dtype=bf); act = torch.empty(M, 4 * C, device=dev, dtype=bf)
        us2 = timed(lambda: K.linear_fwd(xin, wfc, act, bias=bfc, epilogue=K.EPI_GELU, aux_out=pre), args.iters, flush)
        report("f_fc GEMM (bias+GELU, 2 outputs)", us2, (M * C + 2 * M * 4 * C) * 2, f"{fl / us2 / 1e6:.0f} TFLOP/s")
        us2b = timed(lambda: K.linear_fwd(xin, wfc, act, bias=bfc, epilogue=K.EPI_GELU_DG, aux_out=pre), args.iters, flush)
        report("f_fc GEMM (bias+GELU, saves gelu')", us2b, (M * C + 2 * M * 4 * C) * 2, f"{fl / us2b / 1e6:.0f} TFLOP/s")
        res = torch.randn(M, C, device=dev, dtype=bf); wo = torch.randn(C, C, device=dev, dtype=bf) * 0.02
        outo = torch.empty(M, C, device=dev, dtype=bf); bo = torch.zeros(C, device=dev)
        us3 = timed(lambda: K.linear_fwd(xin, wo, outo, bias=bo, epilogue=K.EPI_RESIDUAL, aux_in=res), args.iters, flush)
        report("f_o GEMM (bias+residual)", us3, 3 * M * C * 2, f"{2.0 * M * C * C / us3 / 1e6:.0f} TFLOP/s")
        us4 = timed(lambda: K.linear_fwd(xin, wo, outo, bias=bo), args.iters, flush)
        report("f_o GEMM (bias only)", us4, 2 * M * C * 2, f"{2.0 * M * C * C / us4 / 1e6:.0f} TFLOP/s")
        del xin, wfc, pre, act, res, outo
    if "attn" in ops:
        B = args.batch
        qkv = torch.randn(M, 3 * C, device=dev, dtype=bf)
        o = torch.empty(M, C, device=dev, dtype=bf); lse = torch.empty(B * H * T, device=dev)
        us = timed(lambda: K.attention_fwd(qkv, o, lse, B, T, H), args.iters, flush)
        fl = 4.0 * B * H * T * (T + 128) * 64 / 2          # causal, counted on 128-wide blocks incl. the diagonal
        report("attention fwd", us, 4 * M * C * 2, f"{fl / us / 1e6:.0f} TFLOP/s (causal)")
        do = torch.randn_like(o); dqkv = torch.empty_like(qkv); delta = torch.empty(B * H * T, device=dev)
        dbias = torch.zeros(3 * C, device=dev)
        us = timed(lambda: K.attention_bwd(qkv, o, do, lse, delta, dqkv, B, T, H), args.iters, flush)
        report("attention bwd", us, 9 * M * C * 2, f"{2.5 * fl / us / 1e6:.0f} TFLOP/s (causal)")
        us2 = timed(lambda: K.attention_bwd(qkv, o, do, lse, delta, dqkv, B, T, H, dbias=dbias), args.iters, flush)
        report("attention bwd (+qkv bias grad)", us2, 9 * M * C * 2, f"fused bias grad costs {us2 - us:+.1f} us")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
