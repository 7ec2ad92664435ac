from .model_factory import ModelFactory, get_model, MODEL_REGISTRY
from .gpt2 import GPT2Config, GPT2LMHeadModel
from .resnet import resnet32, resnet50, resnet101
from .vgg import vgg11, vgg13, vgg16

__all__ = ["ModelFactory", "get_model", "MODEL_REGISTRY", "GPT2Config", "GPT2LMHeadModel",
           "resnet32", "resnet50", "resnet101", "vgg11", "vgg13", "vgg16"]
