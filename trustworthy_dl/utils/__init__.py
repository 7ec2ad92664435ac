from .metrics import MetricsCollector
from .data_loader import get_dataloader, SyntheticLoader
from .timing import DeviceTimer, time_steps, max_over_ranks

__all__ = ["MetricsCollector", "get_dataloader", "SyntheticLoader", "DeviceTimer", "time_steps", "max_over_ranks"]
