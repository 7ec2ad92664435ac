from .trust_manager import TrustManager, NodeStatus, TrustScore, NodeMetrics, LogicalClock
from .node_monitor import NodeMonitor
from .distributed_trainer import DistributedTrainer, TrainingConfig, NodeConfig, TrainingState

__all__ = ["TrustManager", "NodeStatus", "TrustScore", "NodeMetrics", "LogicalClock", "NodeMonitor",
           "DistributedTrainer", "TrainingConfig", "NodeConfig", "TrainingState"]
