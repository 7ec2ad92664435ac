"""Packaging (the reference ships this as ``setup_py.py`` against a ``src/`` tree that does not exist).
``python setup.py build_ext --inplace`` (or ``python __graft_entry__.py``) compiles the sm_100a
kernel library in-tree with nvcc; console scripts follow reference setup_py.py:61-66."""
import os
import subprocess
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_ext import build_ext as _build_ext

ROOT = os.path.dirname(os.path.abspath(__file__))


class build_ext(_build_ext):
    def run(self):
        sys.path.insert(0, ROOT)
        import __graft_entry__ as ge
        ge.build(force=False, verbose=True)


setup(
    name="trustworthy-distributed-dl-b200",
    version="0.1.0",
    description="Blackwell-native trustworthy distributed deep learning (trust scoring, attack detection, "
                "trust-weighted collectives) — capabilities of Tanmoy058/Trustworthy-Distributed-Deep-Learning",
    packages=find_packages(include=["trustworthy_dl", "trustworthy_dl.*"]),
    package_data={"trustworthy_dl": ["_C/*.so"]},
    python_requires=">=3.9",
    install_requires=["torch>=2.4", "numpy", "scipy", "scikit-learn", "pandas", "pyyaml"],
    # extras of reference setup_py.py:43-59 (dev / docs / experiments) plus the optional chart backend
    extras_require={"dev": ["pytest", "hypothesis", "pytest-timeout", "black", "isort", "mypy", "flake8"],
                    "docs": ["sphinx", "sphinx-rtd-theme"],
                    "experiments": ["wandb", "tensorboard", "jupyter"],
                    "plots": ["matplotlib"]},
    entry_points={"console_scripts": [
        "trustworthy-dl-train=trustworthy_dl.cli:main",
        "trustworthy-dl-experiment=trustworthy_dl.experiments.runner:main",
        "trustworthy-dl-sweep=trustworthy_dl.experiments.sweep:main",
    ]},
    cmdclass={"build_ext": build_ext},
)
