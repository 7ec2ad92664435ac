from .attack_detection import AttackDetector, AttackType, AttackDetectionResult
from .gradient_verification import GradientVerifier
from .parameter_monitor import ParameterMonitor

__all__ = ["AttackDetector", "AttackType", "AttackDetectionResult", "GradientVerifier", "ParameterMonitor"]
