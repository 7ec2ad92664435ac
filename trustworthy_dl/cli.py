"""``trustworthy-dl-train`` console script (reference setup_py.py:63 names
``trustworthy_dl.cli:main``; the module is not shipped).  Trains with the real trainer from a YAML
file and/or flags; under ``torchrun`` every rank joins the process group first."""
from __future__ import annotations

import argparse
import json
import logging
import os
import sys

from .config import load_training_config
from .core.distributed_trainer import DistributedTrainer, TrainingConfig


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="trustworthy-dl-train", description="Trustworthy distributed training")
    p.add_argument("--config", help="YAML file (README schema)")
    p.add_argument("--model", default=None)
    p.add_argument("--size", default=None)
    p.add_argument("--dataset", default=None)
    p.add_argument("--nodes", type=int, default=None)
    p.add_argument("--epochs", type=int, default=None)
    p.add_argument("--batch-size", type=int, default=None)
    p.add_argument("--lr", type=float, default=None)
    p.add_argument("--parallelism", choices=["model", "pipeline", "data", "tensor"], default=None)
    p.add_argument("--trust-threshold", type=float, default=None)
    p.add_argument("--steps-per-epoch", type=int, default=None)
    p.add_argument("--seq-len", type=int, default=None)
    p.add_argument("--no-attack-detection", action="store_true")
    p.add_argument("--no-gradient-verification", action="store_true")
    p.add_argument("--resume", nargs="?", const="latest", default=None, help="checkpoint path (or latest)")
    p.add_argument("--checkpoint-dir", default=None)
    p.add_argument("--device", default=None)
    return p


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    logging.basicConfig(level=logging.INFO)
    over = {}
    for flag, name in (("model", "model_name"), ("size", "model_size"), ("dataset", "dataset_name"),
                       ("nodes", "num_nodes"), ("epochs", "num_epochs"), ("batch_size", "batch_size"),
                       ("lr", "learning_rate"), ("parallelism", "parallelism"),
                       ("trust_threshold", "trust_threshold"), ("steps_per_epoch", "steps_per_epoch"),
                       ("seq_len", "seq_len"), ("checkpoint_dir", "checkpoint_dir"), ("device", "device")):
        v = getattr(args, flag)
        if v is not None:
            over[name] = v
    if args.no_attack_detection:
        over["attack_detection_enabled"] = False
    if args.no_gradient_verification:
        over["gradient_verification_enabled"] = False
    cfg = load_training_config(args.config, **over) if args.config else TrainingConfig.from_kwargs(**over)
    trainer = DistributedTrainer(cfg)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        trainer.setup_distributed_environment(int(os.environ["RANK"]), world,
                                              master_addr=os.environ.get("MASTER_ADDR", "127.0.0.1"),
                                              master_port=int(os.environ.get("MASTER_PORT", "12355")))
    try:
        if args.resume:
            try:
                trainer.load_checkpoint(None if args.resume == "latest" else args.resume)
            except FileNotFoundError as e:
                if args.resume != "latest":
                    raise
                logging.getLogger(__name__).warning("--resume: %s; starting from scratch", e)
        stats = trainer.train()
        if int(os.environ.get("RANK", "0")) == 0:
            print(json.dumps({k: v for k, v in stats.items() if k != "history"}, default=str))
    finally:
        trainer.cleanup()
    return 0


if __name__ == "__main__":
    sys.exit(main())
