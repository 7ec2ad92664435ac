from .adversarial_attacks import AdversarialAttacker, AttackConfig

__all__ = ["AdversarialAttacker", "AttackConfig"]
