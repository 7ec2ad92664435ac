#!/usr/bin/env python
"""Where a data-parallel step spends its time: forward / backward / (security + all-reduce + AdamW), device-timed with CUDA
events around the three engine phases of a non-graph step, max over ranks, plus the individual calls of the last phase.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/phase_times.py
"""
from __future__ import annotations

import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch
import torch.distributed as dist

from trustworthy_dl.engine.gpt2_engine import GPT2Engine
from trustworthy_dl.models.gpt2 import GPT2_PRESETS, GPT2Config
from trustworthy_dl.ops import nn as K
from trustworthy_dl.ops.stats import tensor_statistics_device


def main() -> int:
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg = GPT2Config(**GPT2_PRESETS["gpt2-medium"])
    eng = GPT2Engine(cfg, batch_size=32, seq_len=1024, lr=5e-5, device=dev, security=True, use_graph=False)
    g = torch.Generator().manual_seed(rank)
    idx = torch.randint(0, cfg.vocab_size, (32, 1024), generator=g).to(dev)
    eng.idx.copy_(idx); eng.tgt.copy_(idx)
    for _ in range(3):
        eng._step_body()
    torch.cuda.synchronize()
    names, evs = [], []

    def mark(name):
        e = torch.cuda.Event(enable_timing=True); e.record(); names.append(name); evs.append(e)

    acc = {}
    reps = 5
    for _ in range(reps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        names.clear(); evs.clear()
        mark("start")
        eng.loss_sum.zero_(); eng.g32.zero_(); mark("zero")
        eng._forward(); mark("forward")
        eng._backward(); mark("backward")
        r = eng.reducer
        tensor_statistics_device(eng.res[eng.L].view(-1), out=eng.act_stats); mark("act_stats")
        K.pack_norm(eng.g32, eng.bucket, r.sumsq, eng.attack_scale); mark("pack_norm")
        r.enqueue_reduce(update_trust=True); mark("trust_allreduce")
        K.adam_tick(eng.hyper, eng.lr, eng.betas[0], eng.betas[1])
        K.adamw(eng.p32, eng.p16, eng.bucket, eng.m, eng.v, eng.n_decay, eng.lr, eng.betas[0], eng.betas[1], eng.eps, eng.wd, 1,
                hyper=eng.hyper); mark("adamw")
        torch.cuda.synchronize()
        for i in range(1, len(evs)):
            acc.setdefault(names[i], []).append(evs[i - 1].elapsed_time(evs[i]))
    out = torch.tensor([sum(v) / len(v) for v in acc.values()], dtype=torch.float64, device=dev)
    lo = out.clone()
    if world > 1:
        dist.all_reduce(out, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"world {world}: phase = max over ranks (min over ranks) in ms, eager launches")
        for n, hi, l in zip(acc.keys(), out.tolist(), lo.tolist()):
            print(f"  {n:16s} {hi:8.3f} ({l:8.3f})")
        print(f"  {'total':16s} {sum(out.tolist()):8.3f}")
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
