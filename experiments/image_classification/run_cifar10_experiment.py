#!/usr/bin/env python
"""README.md:102 — image classification with adversarial attacks (ResNet-32 on CIFAR-10-shaped
synthetic data, 2-stage model parallel, one gradient-poisoning node = BASELINE config 1)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from trustworthy_dl.experiments.runner import main

if __name__ == "__main__":
    cfg = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "configs", "resnet32_cifar_mp2.yaml")
    raise SystemExit(main(["--config", cfg, "--attack", *sys.argv[1:]]))
