"""``get_dataloader(name, split, batch_size)`` — called by the reference runner
(experiment_runner.py:24,100-110) but not shipped.  Batches are ``{'input', 'target'}`` dicts
(distributed_trainer.py:395,398,502-503).

There is no network on the build/bench boxes, so every dataset is *synthetic data of the named
shape* (BASELINE.json north star): ``openwebtext`` -> int64 token streams ``[B, T]`` over GPT-2's
50257-token vocabulary with next-token targets; ``cifar10`` -> ``[B,3,32,32]`` + 10 classes;
``imagenet`` -> ``[B,3,224,224]`` + 1000 classes.  The image sets are class-conditional Gaussians
and the token stream a noisy periodic language, so losses actually go down and attacks have a
visible effect.  Batches can be staged in pinned host memory for the end-to-end H2D path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Iterator, Optional

import torch

__all__ = ["get_dataloader", "SyntheticLoader", "DATASETS"]

DATASETS = {
    "openwebtext": dict(kind="tokens", vocab=50257, seq_len=1024),
    "wikitext": dict(kind="tokens", vocab=50257, seq_len=1024),
    "tiny-tokens": dict(kind="tokens", vocab=512, seq_len=64),
    "cifar10": dict(kind="image", shape=(3, 32, 32), classes=10),
    "cifar-10": dict(kind="image", shape=(3, 32, 32), classes=10),
    "imagenet": dict(kind="image", shape=(3, 224, 224), classes=1000),
}


class SyntheticLoader:
    """Deterministic, re-iterable synthetic dataset."""

    def __init__(self, name: str, split: str = "train", batch_size: int = 32, num_batches: Optional[int] = None,
                 seq_len: Optional[int] = None, vocab: Optional[int] = None, seed: int = 0,
                 pin_memory: bool = False, device: Optional[torch.device] = None, image_size: Optional[int] = None,
                 num_classes: Optional[int] = None):
        key = name.lower()
        if key not in DATASETS:
            raise ValueError(f"unknown dataset {name!r}; have {sorted(DATASETS)}")
        self.spec = dict(DATASETS[key])
        if seq_len:
            self.spec["seq_len"] = seq_len
        if vocab:
            self.spec["vocab"] = vocab
        if image_size and self.spec["kind"] == "image":
            self.spec["shape"] = (3, image_size, image_size)
        if num_classes and self.spec["kind"] == "image":
            self.spec["classes"] = num_classes
        self.name, self.split, self.batch_size = key, split, int(batch_size)
        default = 20 if split == "train" else 4
        self.num_batches = int(num_batches) if num_batches is not None else default
        self.seed = seed + (0 if split == "train" else 1_000_003)
        self.pin_memory = pin_memory and torch.cuda.is_available()
        self.device = device
        if self.spec["kind"] == "image":
            g = torch.Generator().manual_seed(1729)
            c, h, w = self.spec["shape"]
            # low-rank class prototypes: cheap to build even for imagenet-sized images
            self._proto_c = torch.randn(self.spec["classes"], c, 1, 1, generator=g)
            self._proto_s = torch.randn(self.spec["classes"], 1, h // 8 or 1, w // 8 or 1, generator=g)

    def __len__(self) -> int:
        return self.num_batches

    def _make(self, i: int) -> Dict[str, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 7_919 + i)
        B = self.batch_size
        if self.spec["kind"] == "tokens":
            T, V = self.spec["seq_len"], self.spec["vocab"]
            start = torch.randint(0, V, (B, 1), generator=g)
            stride = torch.randint(1, 17, (B, 1), generator=g)
            ramp = torch.arange(T + 1).unsqueeze(0)
            seq = (start + stride * ramp) % V
            noise = torch.rand(B, T + 1, generator=g) < 0.05
            seq = torch.where(noise, torch.randint(0, V, (B, T + 1), generator=g), seq)
            x, y = seq[:, :-1].contiguous(), seq[:, 1:].contiguous()
        else:
            c, h, w = self.spec["shape"]
            y = torch.randint(0, self.spec["classes"], (B,), generator=g)
            spatial = torch.nn.functional.interpolate(self._proto_s[y], size=(h, w), mode="nearest")
            x = 0.7 * (self._proto_c[y] + spatial) + torch.randn(B, c, h, w, generator=g)
        if self.pin_memory:
            x, y = x.pin_memory(), y.pin_memory()
        if self.device is not None:
            x, y = x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)
        return {"input": x, "target": y}

    def __iter__(self) -> Iterator[Dict[str, torch.Tensor]]:
        for i in range(self.num_batches):
            yield self._make(i)


def get_dataloader(name: str, split: str = "train", batch_size: int = 32, **kwargs) -> SyntheticLoader:
    return SyntheticLoader(name, split=split, batch_size=batch_size, **kwargs)
