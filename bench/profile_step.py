#!/usr/bin/env python
"""One GPT-2-medium training step outside a CUDA graph, bracketed by cudaProfilerStart/Stop, for
`ncu --profile-from-start off`.  See profiles/README.md for the command lines."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch

from trustworthy_dl.engine.gpt2_engine import GPT2Engine
from trustworthy_dl.models.gpt2 import GPT2Config, GPT2_PRESETS

p = argparse.ArgumentParser()
p.add_argument("--batch", type=int, default=8)
p.add_argument("--layers", type=int, default=None, help="profile a truncated stack (kernels are per-layer identical)")
args = p.parse_args()
kw = dict(GPT2_PRESETS["gpt2-medium"])
if args.layers:
    kw["n_layer"] = args.layers
cfg = GPT2Config(**kw)
dev = torch.device("cuda", 0)
eng = GPT2Engine(cfg, batch_size=args.batch, seq_len=1024, device=dev, use_graph=False, security=True)
idx = torch.randint(0, cfg.vocab_size, (args.batch, 1024), device=dev)
for _ in range(2):
    eng.train_step(idx, idx)
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng.train_step(idx, idx)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled one step; loss ok")
