#!/usr/bin/env python
"""Where a data-parallel step spends its time, device-timed with CUDA events, max over ranks:

* serial schedule (TDL_OVERLAP=0 semantics): forward / backward / per-segment pack+reduce+AdamW, each reduction with
  its NVLink bytes per direction and the achieved fraction of the 0.9 TB/s link rate;
* overlapped schedule (the default): whole step, and the exposed tail after the last backward kernel.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/phase_times.py [--batch 32]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch
import torch.distributed as dist

from trustworthy_dl.engine.gpt2_engine import GPT2Engine
from trustworthy_dl.models.gpt2 import GPT2_PRESETS, GPT2Config
from trustworthy_dl.ops import nn as K

LINK_BYTES_PER_S = 0.9e12


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--model", default="gpt2-medium")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--zero", type=int, default=0, help="1 = ZeRO-1: sharded AdamW + integrity-checked parameter all-gather (K9)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    cfg = GPT2Config(**GPT2_PRESETS[args.model])
    T = min(1024, cfg.n_positions)
    eng = GPT2Engine(cfg, batch_size=args.batch, seq_len=T, lr=5e-5, device=dev, security=True, use_graph=False, zero=args.zero)
    g = torch.Generator().manual_seed(rank)
    idx = torch.randint(0, cfg.vocab_size, (args.batch, T), generator=g).to(dev)
    eng.idx[0].copy_(idx); eng.tgt[0].copy_(idx)
    for _ in range(3):
        eng._step_body()
    torch.cuda.synchronize()
    names, evs = [], []

    def mark(name):
        e = torch.cuda.Event(enable_timing=True); e.record(); names.append(name); evs.append(e)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    acc = {}
    for _ in range(args.reps):
        sync()
        names.clear(); evs.clear()
        mark("start")
        eng.begin_step(); mark("zero+tick")
        eng._forward(0); mark("forward")
        eng._backward(0, 0); mark("backward")
        for i, seg in enumerate(eng.segments):
            eng._reduce_segment(i, overlapped=False); mark(f"seg{i} pack+reduce+adamw [{seg.name}, {seg.n * 2 / 1e6:.0f} MB bf16]")
        eng.reduce_and_update(overlapped_already=True); mark("finalize")
        torch.cuda.synchronize()
        for i in range(1, len(evs)):
            acc.setdefault(names[i], []).append(evs[i - 1].elapsed_time(evs[i]))
    # reduction alone (K2 only) per segment
    red = {}
    if world > 1:
        r = eng.reducer
        for i, seg in enumerate(eng.segments):
            ts = []
            for _ in range(args.reps):
                K.pack_norm(eng.g32[seg.off:seg.off + seg.n], eng.bucket[seg.off:seg.off + seg.n], r.sumsq, eng.attack_scale)
                sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); r.enqueue_bucket(seg.off, seg.n, owner_only=bool(args.zero)); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            red[i] = min(ts)
        r.enqueue_finalize(update_trust=False)
    # the same bucket through K4 (plain mean all-reduce: multimem through the switch, or P2P) and through NCCL
    cmp_ms = {}
    if world > 1:
        r = eng.reducer

        def timeit(fn, reps=args.reps):
            ts = []
            for _ in range(reps):
                sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            return min(ts)

        def k2_all():
            for seg in eng.segments:
                r.enqueue_bucket(seg.off, seg.n, owner_only=bool(args.zero))
        eng.bucket.normal_()
        cmp_ms["K2 trust-weighted all-reduce + median/cosine statistics (all segments)"] = timeit(k2_all)
        r.enqueue_finalize(update_trust=False)
        cmp_ms["K4 plain mean all-reduce (" + ("multimem.ld_reduce + multimem.st" if r.k4_multimem else "P2P") + ")"] = \
            timeit(lambda: r.enqueue_sum_allreduce(0, eng.P))
        cmp_ms["NCCL all_reduce (bf16, same bucket)"] = timeit(lambda: dist.all_reduce(eng.bucket))
    # overlapped schedule: whole step
    whole = []
    for _ in range(args.reps):
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng._step_body(); e1.record(); torch.cuda.synchronize()
        whole.append(e0.elapsed_time(e1))
    keys = list(acc.keys())
    vals = [sum(v) / len(v) for v in acc.values()] + [red.get(i, 0.0) for i in range(len(eng.segments))] + [sum(whole) / len(whole)]
    out = torch.tensor(vals, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(out, op=dist.ReduceOp.MAX)
    out = out.tolist()
    if rank == 0:
        nk = len(keys)
        serial = sum(out[:nk])
        print(f"world {world}, {args.model}, batch {args.batch} x {T}, zero={args.zero}: ms, max over ranks, eager launches")
        for n, v in zip(keys, out[:nk]):
            print(f"  {n:64s} {v:8.3f}")
        print(f"  {'serial total':64s} {serial:8.3f}")
        print(f"  {'overlapped schedule, whole step':64s} {out[-1]:8.3f}   (hidden: {serial - out[-1]:.3f} ms)")
        if world > 1:
            print("  trust-weighted reduction alone (exchange + coefficients + K2 + barrier); NVLink bytes RECEIVED per rank =")
            print("  load responses for its slice from W-1 peers" + ("" if args.zero else " + the W-1 reduced slices its peers store into it") + ":")
            tot_b, tot_t = 0.0, 0.0
            for i, seg in enumerate(eng.segments):
                b = seg.n * 2 * (world - 1) / world * (1 if args.zero else 2)
                t = out[nk + i] / 1e3
                tot_b += b; tot_t += t
                print(f"    seg{i}: {b / 1e6:8.1f} MB in {t * 1e3:7.3f} ms = {b / t / 1e9:6.1f} GB/s = {b / t / LINK_BYTES_PER_S:.2f} of link rate")
            print(f"    all : {tot_b / 1e6:8.1f} MB in {tot_t * 1e3:7.3f} ms = {tot_b / tot_t / 1e9:6.1f} GB/s = {tot_b / tot_t / LINK_BYTES_PER_S:.2f} of link rate")
        if cmp_ms:
            print(f"  same {eng.P * 2 / 1e6:.0f} MB bf16 bucket, whole (this rank's timings):")
            for k, v in cmp_ms.items():
                print(f"    {k:80s} {v:7.3f} ms")
        print(json.dumps({"world": world, "serial_ms": serial, "overlapped_ms": out[-1], "allreduce_ms": cmp_ms,
                          "phases": dict(zip(keys, out[:nk])), "reduce_ms": out[nk:nk + len(eng.segments)]}))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
