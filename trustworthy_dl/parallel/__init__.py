"""Parallel execution strategies behind :class:`trustworthy_dl.DistributedTrainer`."""
from __future__ import annotations


def make_strategy(trainer):
    mode = trainer.config.parallelism
    if mode == "data":
        from .data_parallel import DataParallelStrategy
        return DataParallelStrategy(trainer)
    if mode in ("model", "pipeline"):
        from .pipeline_fused import FusedPipelineStrategy, fused_pipeline_applicable
        why = fused_pipeline_applicable(trainer)
        if why is None:
            return FusedPipelineStrategy(trainer)          # GPT-2 stages on the sm_100a stage engine
        if trainer.config.engine == "fused":
            raise ValueError(f"engine: fused is not available for this pipeline configuration: {why}")
        from .pipeline import PipelineStrategy
        return PipelineStrategy(trainer)
    if mode == "tensor":
        from .tensor_parallel import TensorParallelStrategy
        return TensorParallelStrategy(trainer)
    raise ValueError(f"unknown parallelism {mode!r}")
