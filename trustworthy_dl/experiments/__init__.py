from .runner import ExperimentConfig, ExperimentRunner, main

__all__ = ["ExperimentConfig", "ExperimentRunner", "main"]
