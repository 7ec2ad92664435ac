"""Parallel execution strategies behind :class:`trustworthy_dl.DistributedTrainer`."""
from __future__ import annotations


def make_strategy(trainer):
    mode = trainer.config.parallelism
    if mode == "data":
        from .data_parallel import DataParallelStrategy
        return DataParallelStrategy(trainer)
    if mode in ("model", "pipeline"):
        from .pipeline import PipelineStrategy
        return PipelineStrategy(trainer)
    if mode == "tensor":
        from .tensor_parallel import TensorParallelStrategy
        return TensorParallelStrategy(trainer)
    raise ValueError(f"unknown parallelism {mode!r}")
