"""Process-group plumbing: NCCL on GPUs, gloo on CPU, rendezvous
on 127.0.0.1; typed p2p send/recv with a small shape header so stages can exchange activations
whose shapes they have never seen; replicated-decision helper (all-gather of small float vectors).
The reference only ever calls ``init_process_group('nccl')`` and ``destroy_process_group``
(distributed_trainer.py:99-114,523-527) and moves no data."""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

__all__ = ["init_process_group", "rank", "world_size", "send_tensor", "recv_tensor", "isend_tensor",
           "all_gather_vector", "broadcast_float", "broadcast_tensor", "is_gloo", "is_short_batch"]

_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32, torch.float64, torch.uint8, torch.bool]


def init_process_group(backend: Optional[str] = None, rank_: Optional[int] = None, world: Optional[int] = None,
                       master_addr: str = "127.0.0.1", master_port: int = 12355) -> Tuple[int, int]:
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    os.environ.setdefault("MASTER_ADDR", master_addr)
    os.environ.setdefault("MASTER_PORT", str(master_port))
    rank_ = int(os.environ.get("RANK", 0)) if rank_ is None else rank_
    world = int(os.environ.get("WORLD_SIZE", 1)) if world is None else world
    backend = backend or ("nccl" if torch.cuda.is_available() else "gloo")
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", rank_ % max(1, torch.cuda.device_count())))
        torch.cuda.set_device(local)
        dist.init_process_group(backend, rank=rank_, world_size=world,
                                device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend, rank=rank_, world_size=world)
    return rank_, world


def rank() -> int:
    return dist.get_rank() if dist.is_initialized() else 0


def world_size() -> int:
    return dist.get_world_size() if dist.is_initialized() else 1


def is_gloo() -> bool:
    return dist.is_initialized() and dist.get_backend() == "gloo"


def _wire_device(device: torch.device) -> torch.device:
    """Device a message travels on: host memory under gloo, always a GPU under NCCL (which has no CPU transport)."""
    if is_gloo():
        return torch.device("cpu")
    device = torch.device(device)
    return device if device.type == "cuda" else torch.device("cuda", torch.cuda.current_device())


def _comm_device(t: torch.Tensor) -> torch.device:
    return _wire_device(t.device)


def _header(t: torch.Tensor) -> torch.Tensor:
    h = torch.full((8,), -1, dtype=torch.int64)
    h[0] = _DTYPES.index(t.dtype)
    h[1] = t.dim()
    for i, s in enumerate(t.shape):
        h[2 + i] = s
    return h


def isend_tensor(t: torch.Tensor, dst: int, tag: int = 0, header: bool = True) -> List:
    """Non-blocking typed send; returns work handles (keep the tensor alive until waited).  ``header=False``: the
    receiver already knows shape and dtype (see :func:`recv_tensor`), only the payload travels."""
    dev = _comm_device(t)
    payload = t.detach().contiguous().to(dev)
    if not header:
        return [dist.isend(payload, dst, tag=tag), payload]
    hdr = _header(t).to(dev)
    return [dist.isend(hdr, dst, tag=tag), dist.isend(payload, dst, tag=tag), payload, hdr]


def send_tensor(t: torch.Tensor, dst: int, tag: int = 0) -> None:
    for w in isend_tensor(t, dst, tag)[:2]:
        w.wait()


def recv_tensor(src: int, device: torch.device, tag: int = 0, like: Optional[Tuple[Tuple[int, ...], torch.dtype]] = None) -> torch.Tensor:
    """Typed receive.  Without ``like`` the shape/dtype header is received first and read on the host (one
    synchronisation); with ``like = (shape, dtype)`` -- a message whose layout both sides already know -- the payload is
    received straight into a new buffer and nothing waits on the host."""
    dev = _wire_device(device)
    if like is not None:
        buf = torch.empty(like[0], dtype=like[1], device=dev)
        dist.recv(buf, src, tag=tag)
        return buf.to(device)
    hdr = torch.empty(8, dtype=torch.int64, device=dev)
    dist.recv(hdr, src, tag=tag)
    h = hdr.cpu().tolist()
    shape = [int(s) for s in h[2:2 + int(h[1])]]
    buf = torch.empty(shape, dtype=_DTYPES[int(h[0])], device=dev)
    dist.recv(buf, src, tag=tag)
    return buf.to(device)


def all_gather_vector(v: torch.Tensor) -> torch.Tensor:
    """[W, len(v)] float64 gather of a small per-rank vector (decision inputs)."""
    W = world_size()
    if W == 1:
        return v.reshape(1, -1).double().cpu()
    dev = torch.device("cpu") if is_gloo() else torch.device("cuda", torch.cuda.current_device())
    x = v.detach().double().reshape(-1).to(dev)
    out = torch.empty(W * x.numel(), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(out, x)
    return out.reshape(W, -1).cpu()


_short_batch_warned = False


def is_short_batch(batch, expected: int) -> bool:
    """The fused engines run static shapes (captured graphs, fixed activation buffers): a batch that is not the
    configured size -- the last one of a loader without ``drop_last`` -- is skipped by every rank, with one log line."""
    global _short_batch_warned
    n = int(batch["input"].shape[0])
    if n == expected:
        return False
    if not _short_batch_warned:
        import logging
        logging.getLogger(__name__).warning("skipping a batch of %d samples (configured batch_size %d): the fused engines run "
                                            "static shapes; use drop_last or a divisible dataset", n, expected)
        _short_batch_warned = True
    return True


def broadcast_tensor(t: torch.Tensor, src: int) -> None:
    """In-place broadcast of ``t`` from rank ``src`` (staged through the host on the gloo backend)."""
    if world_size() == 1:
        return
    if is_gloo() and t.is_cuda:
        h = t.detach().cpu()
        dist.broadcast(h, src)
        t.copy_(h)
    else:
        dist.broadcast(t.detach(), src)


def broadcast_float(value: float, src: int) -> float:
    if world_size() == 1:
        return value
    dev = torch.device("cpu") if is_gloo() else torch.device("cuda", torch.cuda.current_device())
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.broadcast(t, src)
    return float(t.item())
