"""trustworthy_dl — Blackwell-native trustworthy distributed training.

Public API of the reference (README.md:59-61): ``from trustworthy_dl import DistributedTrainer,
TrustManager``; ``trustworthy_dl.models.get_model``; ``trustworthy_dl.attacks.AdversarialAttacker``.
"""
from .core.trust_manager import TrustManager, NodeStatus, TrustScore, NodeMetrics
from .core.distributed_trainer import DistributedTrainer, TrainingConfig, NodeConfig, TrainingState
from .security.attack_detection import AttackDetector, AttackType, AttackDetectionResult
from .security.gradient_verification import GradientVerifier
from .security.parameter_monitor import ParameterMonitor
from .attacks.adversarial_attacks import AdversarialAttacker, AttackConfig
from .models.model_factory import ModelFactory, get_model
from .utils.metrics import MetricsCollector
from .utils.data_loader import get_dataloader

__version__ = "0.1.0"

__all__ = ["TrustManager", "NodeStatus", "TrustScore", "NodeMetrics", "DistributedTrainer", "TrainingConfig",
           "NodeConfig", "TrainingState", "AttackDetector", "AttackType", "AttackDetectionResult",
           "GradientVerifier", "ParameterMonitor", "AdversarialAttacker", "AttackConfig", "ModelFactory",
           "get_model", "MetricsCollector", "get_dataloader", "__version__"]
