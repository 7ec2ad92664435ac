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


_CTX = None


def _context():
    """Fork-server start method with torch and the package preloaded: every worker is forked from one clean, single-threaded
    server process that has already paid for ``import torch`` (several seconds per process under ``spawn``)."""
    global _CTX
    if _CTX is None:
        _CTX = mp.get_context("forkserver")
        _CTX.set_forkserver_preload(["torch", "torch.distributed", "numpy", "trustworthy_dl", "trustworthy_dl.parallel"])
    return _CTX


def run_distributed(fn, world: int, *args, timeout: float = 240.0):
    ctx = _context()
    mgr = ctx.Manager()
    results = mgr.dict()
    port = free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, results)) for r in range(world)]
    for p in procs:
        p.start()
    import time
    deadline = time.time() + timeout
    while any(p.is_alive() for p in procs) and time.time() < deadline:
        seen = dict(results)
        if any(v[0] == "err" for v in seen.values()):
            break                                  # a rank failed: its peers would wait for it in a collective until the deadline
        if any((not p.is_alive()) and i not in seen for i, p in enumerate(procs)):
            time.sleep(0.5)                        # (the result may still be in flight)
            if any((not p.is_alive()) and i not in dict(results) for i, p in enumerate(procs)):
                break                              # a rank died without reporting: same thing
        time.sleep(0.1)
    alive = [i for i, p in enumerate(procs) if p.is_alive()]
    if alive:
        if any(v[0] == "err" for v in dict(results).values()):
            time.sleep(1.0)                        # let the other ranks report their own exception, if any
        failed = {r: v[1] for r, v in dict(results).items() if v[0] == "err"}
        for i, p in enumerate(procs):
            if not p.is_alive() and i not in dict(results) and i not in failed:
                failed[i] = f"process exited with code {p.exitcode} without reporting a result"
        for p in procs:
            if p.is_alive():
                p.kill()
        assert not failed, "\n".join(f"rank {r} failed:\n{tb}" for r, tb in sorted(failed.items()))
        raise AssertionError(f"distributed test timed out (ranks {alive} still running after {timeout:.0f} s)")
    out = []
    for r in range(world):
        status, val = results.get(r, ("err", "no result"))
        assert status == "ok", f"rank {r} failed:\n{val}"
        out.append(val)
    return out
