"""Python entry points of the sm_100a kernels (+ the CPU restatements used as oracles)."""
from . import reference_impl
from .stats import tensor_statistics, mean_std, STAT_NAMES

__all__ = ["reference_impl", "tensor_statistics", "mean_std", "STAT_NAMES"]
