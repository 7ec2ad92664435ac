#!/usr/bin/env python
"""Stock-library comparison arm: the same GPT-2-medium training step on torch eager +
cuBLAS/SDPA + fused AdamW (+ NCCL DDP for N>1), optionally with the *reference-style* security path —
the unmodified reference ``attack_detector.py`` / ``trust_manager.py`` (copied verbatim into the
git-ignored ``baseline/_ref``) fed the way the reference trainer feeds them: a full activation D2H copy
and numpy/scipy statistics on the host every step (distributed_trainer.py:167-170, attack_detector.py:78).

This is NOT the ``--impl reference`` arm (the reference's trainer cannot be imported); it is the bar the
north star names: reference semantics on stock libraries, same box, same model/config.
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _load_ref(name):
    for d in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        path = os.path.join(d, f"{name}.py")
        if os.path.exists(path):
            spec = importlib.util.spec_from_file_location(f"_ref_{name}", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    return None


def main(argv=None) -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--steps", type=int, default=10)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--batch", type=int, default=32)
    p.add_argument("--seq-len", type=int, default=1024)
    p.add_argument("--model", default="gpt2-medium")
    p.add_argument("--security", choices=["off", "reference"], default="off")
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--compile", action="store_true", help="wrap the stock model in torch.compile (baseline arm only)")
    args = p.parse_args(argv)
    import torch
    import torch.distributed as dist
    from trustworthy_dl.models.gpt2 import GPT2Config, GPT2LMHeadModel, GPT2_PRESETS

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)
    cfg = GPT2Config(**GPT2_PRESETS[args.model])
    model = GPT2LMHeadModel(cfg).to(dev)
    if args.compile:
        model = torch.compile(model)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
    opt = torch.optim.AdamW(model.parameters(), lr=5e-5, weight_decay=0.01, fused=True)
    core = model.module if world > 1 else model
    core = getattr(core, "_orig_mod", core)
    detector = trust = None
    if args.security == "reference":
        det_mod, tm_mod = _load_ref("attack_detector"), _load_ref("trust_manager")
        if det_mod is None or tm_mod is None:
            print(json.dumps({"impl": "torch_baseline", "unavailable": "reference modules not found"}))
            return 0
        detector, trust = det_mod.AttackDetector(), tm_mod.TrustManager(world)
    B, T = args.batch, args.seq_len
    g = torch.Generator().manual_seed(1234 + rank)
    data = [(torch.randint(0, cfg.vocab_size, (B, T), generator=g).to(dev), torch.randint(0, cfg.vocab_size, (B, T), generator=g).to(dev))
            for _ in range(4)]
    hidden = {}
    core.transformer.h[-1].register_forward_hook(lambda m, i, o: hidden.__setitem__("x", o.detach()))

    def step(i):
        idx, tgt = data[i % 4]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = model(idx)
            # bf16 logits straight into the loss (no fp32 copy of the 3.3 GB tensor): the arm is not handicapped
            loss = torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), tgt.reshape(-1))
        loss.backward()
        if detector is not None:                       # reference-style: whole activation to the host, numpy/scipy stats
            flagged = detector.detect_output_anomaly(hidden["x"].float(), rank, i)
            gn = torch.stack([p.grad.norm() for p in core.parameters() if p.grad is not None])
            trust.update_trust_score(rank, 1.0 if flagged else 0.0, 1.0 if bool(torch.isfinite(gn).all()) else 0.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        return loss

    for i in range(max(3, args.warmup)):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(i)
    e1.record(); e1.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        print(json.dumps({"impl": "torch_baseline", "security": args.security, "metric": "gpt2_medium_train_tokens_per_s",
                          "value": B * T * world / (ms / args.steps / 1e3), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
                          "ms_per_step": ms / args.steps, "dtype": "bf16", "data": "synthetic", "final_loss": float(loss.item()),
                          "config": {"model": args.model, "global_batch": B * world, "seq_len": T, "parallelism": f"ddp{world}"}}))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
