#!/usr/bin/env python
"""Device-timed microbenchmark of the tcgen05 GEMM on the GPT-2-medium shapes vs torch.matmul (cuBLAS),
with an L2 flush between timed iterations.  Prints TFLOP/s and fraction of the measured bf16 peak."""
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch

from trustworthy_dl.ops import nn as K

dev = torch.device("cuda", 0)
peak = 1661.4
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    pass
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
M, C, V = 32768, 1024, 50304
shapes = [("f_qkv", "nt", M, 3 * C, C), ("f_o", "nt", M, C, C), ("f_fc", "nt", M, 4 * C, C), ("f_proj", "nt", M, C, 4 * C),
          ("d_proj", "dgrad", M, 4 * C, C), ("d_fc", "dgrad", M, C, 4 * C), ("w_fc", "wgrad", 4 * C, C, M), ("w_o", "wgrad", C, C, M),
          ("f_head", "nt", M, V, C)]
block_n = int(os.environ.get("BLOCK_N", "0"))


def timeit(fn, iters=8):
    fn(); fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


for name, kind, m, n, k in shapes:
    if kind == "nt":
        a = torch.randn(m, k, device=dev).bfloat16(); b = torch.randn(n, k, device=dev).bfloat16()
        out = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        ours = lambda: K.gemm(a, b, out, m, n, k, lda=k, ldb=k, ldc=n, block_n=block_n)
        ref = lambda: torch.matmul(a, b.t(), out=out)
    elif kind == "dgrad":
        a = torch.randn(m, k, device=dev).bfloat16(); b = torch.randn(k, n, device=dev).bfloat16()
        out = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        ours = lambda: K.gemm(a, b, out, m, n, k, lda=k, ldb=n, ldc=n, b_mn=True, block_n=block_n)
        ref = lambda: torch.matmul(a, b, out=out)
    else:
        a = torch.randn(k, m, device=dev).bfloat16(); b = torch.randn(k, n, device=dev).bfloat16()
        out32 = torch.zeros(m, n, device=dev)
        sk = 0                                   # library picks the split
        ours = lambda: K.gemm(a, b, out32, m, n, k, lda=m, ldb=n, ldc=n, a_mn=True, b_mn=True, atomic_f32=True, split_k=sk)
        outb = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
        ref = lambda: torch.matmul(a.t(), b, out=outb)
    t_o, t_r = timeit(ours), timeit(ref)
    fl = 2.0 * m * n * k
    print(f"{name:7s} {kind:6s} M={m:6d} N={n:6d} K={k:6d}  ours {t_o:8.1f} us {fl / t_o / 1e6:7.1f} TF ({fl / t_o / 1e6 / peak:4.2f} of measured)"
          f"   cuBLAS {t_r:8.1f} us {fl / t_r / 1e6:7.1f} TF", flush=True)
