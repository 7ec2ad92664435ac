#!/usr/bin/env python
"""NVLink peer-memory probe on symmetric memory: (a) pull = read the peer's buffer into a local one with a torch copy
kernel, (b) push = write a local buffer into the peer's, (c) the trust all-reduce kernel alone on the same amount of data.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 bench/p2p_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
import torch.distributed as dist

from trustworthy_dl.parallel.symm import SymmetricBuffer, SymmetricTrustReducer


def timed(fn, iters=10):
    fn(); torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    n = 355_000_000 // 8 * 8
    buf = SymmetricBuffer(n, torch.bfloat16, dev)
    peer = (rank + 1) % world
    import torch.distributed._symmetric_memory as symm_mem
    peer_t = buf.handle.get_buffer(peer, (n,), torch.bfloat16)
    loc = torch.empty(n, dtype=torch.bfloat16, device=dev)
    half = n // world
    ms = timed(lambda: loc[:half].copy_(peer_t[:half]))
    if rank == 0:
        print(f"pull  {half * 2 / 1e6:.0f} MB from peer: {ms:.3f} ms = {half * 2 / ms / 1e6:.0f} GB/s per direction (all ranks at once)")
    ms = timed(lambda: peer_t[:half].copy_(loc[:half]))
    if rank == 0:
        print(f"push  {half * 2 / 1e6:.0f} MB to peer:   {ms:.3f} ms = {half * 2 / ms / 1e6:.0f} GB/s per direction")
    red = SymmetricTrustReducer(n, torch.bfloat16, dev)
    red.buffer.normal_()
    red.enqueue_local_norm()
    ms = timed(lambda: red.enqueue_reduce(update_trust=False))
    if rank == 0:
        print(f"trust all-reduce of {n * 2 / 1e6:.0f} MB (exchange + coefficients + reduce + exchange + finalize): {ms:.3f} ms "
              f"-> {n * 2 * (world - 1) / world / ms / 1e6:.0f} GB/s read + the same written per rank")
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
