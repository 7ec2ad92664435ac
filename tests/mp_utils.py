"""Spawn helper for world_size>1 CPU tests (gloo, 127.0.0.1)."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, results):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(1)
    try:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        results[rank] = ("ok", fn(rank, world, *args))
    except Exception:
        results[rank] = ("err", traceback.format_exc())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_distributed(fn, world: int, *args, timeout: float = 240.0):
    ctx = mp.get_context("spawn")
    mgr = ctx.Manager()
    results = mgr.dict()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, results)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout)
    alive = [p for p in procs if p.is_alive()]
    for p in alive:
        p.kill()
    assert not alive, "distributed test timed out"
    out = []
    for r in range(world):
        status, val = results.get(r, ("err", "no result"))
        assert status == "ok", f"rank {r} failed:\n{val}"
        out.append(val)
    return out
