#!/usr/bin/env python
"""Device-timed runs of the other BASELINE.json configurations through the public API
(`DistributedTrainer` + strategy), one JSON line per run:

  vgg16_mp8     VGG-16 224x224, 8-stage pipeline (1F1B, NCCL p2p), trust scoring + gradient verification
  resnet50_dp   ResNet-50, data parallel with the fused trust-weighted all-reduce (symmetric memory)
  gpt2_tp8      GPT-2 medium seq 1024 bf16, tensor parallel (fused all-gather->GEMM / GEMM->reduce-scatter),
                one rank injects poisoned gradients
  gpt2_pp4dp2   GPT-2 medium pipeline x4 x data x2, rank compromised mid-run -> trust drop -> reassignment

Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/config_bench.py --config NAME
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch
import torch.distributed as dist

from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
from trustworthy_dl.utils.data_loader import get_dataloader


def build(name: str, world: int, args):
    att = None
    base = args.impl == "baseline"
    sec = dict(attack_detection_enabled=not base, gradient_verification_enabled=not base)
    eng = "eager" if base else "auto"
    if name == "vgg16_mp8":
        cfg = TrainingConfig(model_name="vgg16", dataset_name="imagenet", batch_size=args.batch or 64, num_nodes=world,
                             learning_rate=1e-3, checkpoint_interval=0, parallelism="pipeline", micro_batches=8, dtype="bf16",
                             optimizer="sgd", **sec)
        data = ("imagenet", {})
        unit = "images/s"
    elif name == "resnet50_dp":
        cfg = TrainingConfig(model_name="resnet50", dataset_name="imagenet", batch_size=(args.batch or 32) * world, num_nodes=world,
                             learning_rate=1e-3, checkpoint_interval=0, parallelism="data", dtype="bf16", optimizer="sgd",
                             engine="eager", **sec)
        data = ("imagenet", {})
        unit = "images/s"
    elif name == "gpt2_tp8":
        cfg = TrainingConfig(model_name="gpt2", model_size="medium", dataset_name="openwebtext", batch_size=args.batch or 32,
                             num_nodes=world, learning_rate=5e-5, checkpoint_interval=0, parallelism="tensor", tp_size=world,
                             dtype="bf16", seq_len=1024, optimizer="adamw", engine=eng, **sec)
        if base:
            os.environ["TDL_TP_FUSED"] = "0"          # Megatron-style arm: torch matmul + NCCL all-gather / reduce-scatter
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[min(3, world - 1)], intensity=0.5,
                                               start_step=args.attack_step))
        data = ("openwebtext", {"seq_len": 1024})
        unit = "tokens/s"
    elif name == "gpt2_pp4dp2":
        dp = args.dp if args.dp else (2 if world % 2 == 0 and world >= 4 else 1)
        cfg = TrainingConfig(model_name="gpt2", model_size="medium", dataset_name="openwebtext", batch_size=args.batch or 128 * dp,
                             num_nodes=world, learning_rate=5e-5, checkpoint_interval=0, parallelism="pipeline", pp_size=world // dp,
                             dp_size=dp, micro_batches=args.micro, dtype="bf16", seq_len=1024, optimizer="adamw", migration_model="nvlink",
                             engine=eng, pp_virtual=args.pp_virtual, **sec)
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[min(2, world - 1)], intensity=0.5,
                                               start_step=args.attack_step))
        data = ("openwebtext", {"seq_len": 1024})
        unit = "tokens/s"
    else:
        raise SystemExit(f"unknown config {name}")
    if base or args.no_attack:
        att = None
    if att is not None:
        att.activate_attacks()
    return cfg, att, data, unit


def ddp_baseline(cfg, dname, dkw, unit, args, world, rank, local, dev) -> int:
    """Baseline A of the data-parallel configuration: the same nn.Module under torch DDP (NCCL all-reduce), stock optimizer,
    bf16 autocast, no security."""
    from trustworthy_dl.models.model_factory import ModelFactory
    torch.manual_seed(0)
    model = ModelFactory().create_model(cfg.model_name, cfg.model_size).to(dev)
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
    opt = torch.optim.SGD(model.parameters(), lr=cfg.learning_rate, momentum=0.9, weight_decay=cfg.weight_decay)
    loader = list(get_dataloader(dname, "train", cfg.batch_size, num_batches=4, **dkw))
    core = model.module if world > 1 else model

    def step(i):
        b = loader[i % len(loader)]
        x = b["input"].chunk(world)[rank].to(dev, non_blocking=True)
        y = b["target"].chunk(world)[rank].to(dev, non_blocking=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
        loss = core.loss(out, y)
        loss.backward()
        opt.step(); opt.zero_grad(set_to_none=True)
        return loss

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(i)
    e1.record(); e1.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        total = float(ms.item())
        print(json.dumps({"config": args.config, "impl": "baseline", "strategy": "torch DDP + NCCL all-reduce", "n_gpus": world,
                          "steps": args.steps, "warmup": args.warmup, "value": cfg.batch_size * args.steps / (total / 1e3), "unit": unit,
                          "ms_per_step": total / args.steps, "dtype": cfg.dtype, "data": "synthetic", "global_batch": cfg.batch_size,
                          "final_loss": float(loss.item())}))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


def main() -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--config", required=True)
    p.add_argument("--steps", type=int, default=12)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--batch", type=int, default=0)
    p.add_argument("--attack-step", type=int, default=8)
    p.add_argument("--micro", type=int, default=8, help="micro-batches per step of the pipeline configurations")
    p.add_argument("--no-attack", action="store_true")
    p.add_argument("--dp", type=int, default=0, help="data-parallel replicas of the pipeline configuration (default: 2 from 4 GPUs up)")
    p.add_argument("--pp-virtual", type=int, default=1, help="stages per pipeline rank (interleaved)")
    p.add_argument("--impl", default="ours", choices=["ours", "baseline"],
                   help="baseline = BASELINE.md baseline A: stock kernels (torch eager, cuBLAS/cuDNN/SDPA) + NCCL collectives, "
                        "security off, same box, same model/config")
    args = p.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg, att, (dname, dkw), unit = build(args.config, world, args)
    if args.impl == "baseline" and args.config == "resnet50_dp":
        return ddp_baseline(cfg, dname, dkw, unit, args, world, rank, local, dev)
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    t.device = dev
    n_batches = args.steps + args.warmup
    loader = list(get_dataloader(dname, "train", cfg.batch_size, num_batches=min(n_batches, 4), **dkw))
    per_step = cfg.batch_size * (dkw.get("seq_len", 1))

    def step(i):
        return t.train_step(loader[i % len(loader)])

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    step_ms = []
    losses = []
    for i in range(args.steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        losses.append(step(args.warmup + i))
        e1.record(); e1.synchronize()
        step_ms.append(e0.elapsed_time(e1))
    torch.cuda.synchronize()
    ms = torch.tensor(step_ms, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = ms.cpu().tolist()
    if rank == 0:
        total = sum(ms)
        attack_at = None
        if t.attack_history:
            attack_at = t.attack_history[0]["step"] - args.warmup - 1       # index into the timed steps
        out = {"config": args.config, "impl": args.impl, "strategy": type(t._backend_impl).__name__,
               "fused_engine": bool(getattr(t._backend_impl, "engine", None) is not None or getattr(t._backend_impl, "engines", None)),
               "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "value": per_step * args.steps / (total / 1e3), "unit": unit, "ms_per_step": total / args.steps,
               "ms_per_step_median": sorted(ms)[len(ms) // 2], "step_ms": [round(x, 2) for x in ms], "dtype": cfg.dtype,
               "data": "synthetic", "global_batch": cfg.batch_size, "parallelism": cfg.parallelism,
               "security": {"attack_detection": cfg.attack_detection_enabled, "gradient_verification": cfg.gradient_verification_enabled},
               "first_loss": losses[0], "final_loss": losses[-1],
               "attacks": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history][:4],
               "reassignments": [{k: r.get(k) for k in ("from_node", "to_node", "step", "stages", "migration_time", "measured_time")}
                                 for r in t.reassignment_history][:4],
               "trust": [round(x, 3) for x in t.trust_manager.trust_vector().tolist()],
               "stage_owner": list(t.stage_map.owner) if t.stage_map is not None else None}
        if attack_at is not None and 0 <= attack_at < len(ms):
            out["recovery"] = {"attack_step_ms": ms[attack_at], "steps_before_ms": sum(ms[:attack_at]) / max(1, attack_at),
                               "steps_after_ms": sum(ms[attack_at + 1:]) / max(1, len(ms) - attack_at - 1)}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
