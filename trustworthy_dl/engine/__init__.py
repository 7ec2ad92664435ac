from .gpt2_engine import GPT2Engine

__all__ = ["GPT2Engine"]
