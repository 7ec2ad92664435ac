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
    if name == "vgg16_mp8":
        cfg = TrainingConfig(model_name="vgg16", dataset_name="imagenet", batch_size=args.batch or 64, num_nodes=world,
                             learning_rate=1e-3, checkpoint_interval=0, parallelism="pipeline", micro_batches=8, dtype="bf16",
                             optimizer="sgd")
        data = ("imagenet", {})
        unit = "images/s"
    elif name == "resnet50_dp":
        cfg = TrainingConfig(model_name="resnet50", dataset_name="imagenet", batch_size=(args.batch or 32) * world, num_nodes=world,
                             learning_rate=1e-3, checkpoint_interval=0, parallelism="data", dtype="bf16", optimizer="sgd",
                             engine="eager")
        data = ("imagenet", {})
        unit = "images/s"
    elif name == "gpt2_tp8":
        cfg = TrainingConfig(model_name="gpt2", model_size="medium", dataset_name="openwebtext", batch_size=args.batch or 8,
                             num_nodes=world, learning_rate=5e-5, checkpoint_interval=0, parallelism="tensor", tp_size=world,
                             dtype="bf16", seq_len=1024, optimizer="sgd")
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[min(3, world - 1)], intensity=0.5,
                                               start_step=args.attack_step))
        data = ("openwebtext", {"seq_len": 1024})
        unit = "tokens/s"
    elif name == "gpt2_pp4dp2":
        dp = 2 if world % 2 == 0 and world >= 4 else 1
        cfg = TrainingConfig(model_name="gpt2", model_size="medium", dataset_name="openwebtext", batch_size=args.batch or 32,
                             num_nodes=world, learning_rate=5e-5, checkpoint_interval=0, parallelism="pipeline", pp_size=world // dp,
                             dp_size=dp, micro_batches=8, dtype="bf16", seq_len=1024, optimizer="sgd", migration_model="nvlink")
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[min(2, world - 1)], intensity=0.5,
                                               start_step=args.attack_step))
        data = ("openwebtext", {"seq_len": 1024})
        unit = "tokens/s"
    else:
        raise SystemExit(f"unknown config {name}")
    if att is not None:
        att.activate_attacks()
    return cfg, att, data, unit


def main() -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--config", required=True)
    p.add_argument("--steps", type=int, default=12)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--batch", type=int, default=0)
    p.add_argument("--attack-step", type=int, default=8)
    args = p.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg, att, (dname, dkw), unit = build(args.config, world, args)
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
        out = {"config": args.config, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
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
