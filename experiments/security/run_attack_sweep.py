#!/usr/bin/env python
"""README.md:146-149 — detection rate / false positives / detection latency across attack types and intensities
(ResNet-32 on CIFAR-10-shaped synthetic data, 2-stage model parallel by default; every number is measured)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from trustworthy_dl.experiments.sweep import main

if __name__ == "__main__":
    raise SystemExit(main(sys.argv[1:]))
