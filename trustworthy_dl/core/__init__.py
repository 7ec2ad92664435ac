from .trust_manager import TrustManager, NodeStatus, TrustScore, NodeMetrics, LogicalClock
from .node_monitor import HeartbeatMonitor, NodeMonitor
from .distributed_trainer import DistributedTrainer, TrainingConfig, NodeConfig, TrainingState

__all__ = ["TrustManager", "NodeStatus", "TrustScore", "NodeMetrics", "LogicalClock", "NodeMonitor", "HeartbeatMonitor",
           "DistributedTrainer", "TrainingConfig", "NodeConfig", "TrainingState"]
