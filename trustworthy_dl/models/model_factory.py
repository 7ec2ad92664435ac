"""``ModelFactory`` / ``get_model`` — the reference imports both but ships neither
(distributed_trainer.py:24,118-119; README.md:60).  Names follow README.md:85-92:
``gpt2`` (+ ``-medium``/``-large`` or ``size=``), ``vgg11/13/16``, ``resnet32/50/101``.
``-cifar`` / ``-tiny`` suffixes select reduced-size variants for tests."""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch.nn as nn

from .gpt2 import GPT2Config, GPT2LMHeadModel, GPT2_PRESETS
from .resnet import resnet32, resnet50, resnet101
from .vgg import VGG

__all__ = ["ModelFactory", "get_model", "MODEL_REGISTRY"]


def _gpt2(name: str, size: Optional[str] = None, **kw) -> nn.Module:
    key = name if size in (None, "", "small") else f"gpt2-{size}"
    if key not in GPT2_PRESETS:
        raise ValueError(f"unknown GPT-2 variant {key!r}; have {sorted(GPT2_PRESETS)}")
    return GPT2LMHeadModel(GPT2Config(**{**GPT2_PRESETS[key], **kw}))


def _vgg(depth: int):
    def make(name: str, size: Optional[str] = None, **kw):
        if name.endswith("-cifar") or size == "cifar":
            kw = {"num_classes": 10, "input_size": 32, **kw}
        return VGG(depth, **kw)
    return make


MODEL_REGISTRY: Dict[str, Callable[..., nn.Module]] = {
    **{k: _gpt2 for k in GPT2_PRESETS},
    "resnet32": lambda name, size=None, **kw: resnet32(**kw),
    "resnet50": lambda name, size=None, **kw: resnet50(**kw),
    "resnet101": lambda name, size=None, **kw: resnet101(**kw),
    "vgg11": _vgg(11), "vgg13": _vgg(13), "vgg16": _vgg(16),
    "vgg11-cifar": _vgg(11), "vgg13-cifar": _vgg(13), "vgg16-cifar": _vgg(16),
}


class ModelFactory:
    def available_models(self):
        return sorted(MODEL_REGISTRY)

    def create_model(self, model_name: str, size: Optional[str] = None, **kwargs) -> nn.Module:
        key = model_name.lower().replace("_", "-")
        if key not in MODEL_REGISTRY:
            raise ValueError(f"unknown model {model_name!r}; have {self.available_models()}")
        return MODEL_REGISTRY[key](key, size, **kwargs)


def get_model(name: str, size: Optional[str] = None, **kwargs) -> nn.Module:
    return ModelFactory().create_model(name, size, **kwargs)
