#!/usr/bin/env python
"""K7 (GEMM -> reduce-scatter) and K8 (all-gather -> GEMM) at the shapes of GPT-2 medium under tensor parallelism,
device-timed (CUDA events, max over ranks), each reported as achieved fraction of its roofline
max(FLOPs / measured sustained bf16 peak, NVLink bytes per direction / 0.9 TB/s)  (BASELINE.json north star).

  python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 bench/tp_collective_bench.py [--batch 32]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

import torch
import torch.distributed as dist

from trustworthy_dl.engine.gpt2_tp_engine import GPT2TPEngine
from trustworthy_dl.models.gpt2 import GPT2_PRESETS, GPT2Config
from trustworthy_dl.ops import nn as K, tp_ops


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--step", action="store_true", help="also time whole training steps of the 24-layer TP engine (one CUDA graph)")
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    peaks = {"bf16_tflops_sustained": 1363.3}
    try:
        peaks.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
    except Exception:
        pass
    peak = float(peaks["bf16_tflops_sustained"]) * 1e12
    link = 0.9e12
    cfg = GPT2Config(**{**GPT2_PRESETS["gpt2-medium"], "n_layer": 1})
    e = GPT2TPEngine(cfg, batch_size=args.batch, seq_len=1024, device=dev, use_graph=False, security=False)
    M, Ms, C, Cl, Fl, W = e.M, e.Ms, e.C, e.Cl, e.Fl, world
    e.sym_ln[0][0].normal_(); e.sym_ln[0][1].normal_(); e.sym_dx[0].normal_()
    e.atty[0].normal_(); e.fcg[0].normal_(); e.dfch.normal_(); e.dqkv.normal_(); e.fch[0].fill_(1.0)
    p = "h0."
    mode = {"pull": "P2P pull into the local operand, then the tcgen05 GEMM on local memory (operand kept for the wgrad)",
            "tma": "TMA reads of the peer shards"}[e.k8_mode]
    ag_bytes = (W - 1) * Ms * C * 2            # what one rank pulls (K8) or pushes (K7), one direction
    cases = [
        ("K8 fwd  qkv   AG(ln1)[M,C] @ Wqkv_r[3C/W,C]^T + b", lambda: e._col_parallel(e._off_ln[0][0], e.w16(p + "qkv_w"), e.qkv[0], full=e._lnf(0, 0), bias=e.w32(p + "qkv_b"), block_n=256 if 3 * Cl > 128 else 0), 2.0 * M * 3 * Cl * C, ag_bytes),
        ("K8 fwd  fc    AG(ln2) @ Wfc_r[4C/W,C]^T, GELU+GELU' epilogue", lambda: e._col_parallel(e._off_ln[0][1], e.w16(p + "fc_w"), e.fcg[0], full=e._lnf(0, 1), bias=e.w32(p + "fc_b"), epilogue=K.EPI_GELU_DG, aux_out=e.fch[0]), 2.0 * M * Fl * C, ag_bytes),
        ("K8 bwd  dfc   AG(dx) @ Wproj_r[C,4C/W] * gelu', bias grad", lambda: e._col_parallel(e._off_dx[0], e.w16(p + "proj_w"), e.dfch, full=e.full_a, b_mn=True, epilogue=K.EPI_MUL, aux_in=e.fch[0], colsum=e.grad32(p + "fc_b")), 2.0 * M * Fl * C, ag_bytes),
        ("K8 bwd  datty AG(dx) @ Wo_r[C,C/W]", lambda: e._col_parallel(e._off_dx[0], e.w16(p + "o_w"), e.datty, full=e.full_b, b_mn=True), 2.0 * M * Cl * C, ag_bytes),
        ("K7 fwd  o     atty_r[M,C/W] @ Wo_r^T -> RS + bias + residual", lambda: e._row_parallel(e.atty[0], e.w16(p + "o_w"), e.x2[0], bias=e.w32(p + "o_b"), residual=e.res[0]), 2.0 * M * C * Cl, ag_bytes),
        ("K7 fwd  proj  h_r[M,4C/W] @ Wproj_r^T -> RS + bias + residual", lambda: e._row_parallel(e.fcg[0], e.w16(p + "proj_w"), e.res[1], bias=e.w32(p + "proj_b"), residual=e.x2[0]), 2.0 * M * C * Fl, ag_bytes),
        ("K7 bwd  dln2  dfc_r[M,4C/W] @ Wfc_r -> RS", lambda: e._row_parallel(e.dfch, e.w16(p + "fc_w"), e.dln_s, b_mn=True), 2.0 * M * C * Fl, ag_bytes),
        ("K7 bwd  dln1  dqkv_r[M,3C/W] @ Wqkv_r -> RS", lambda: e._row_parallel(e.dqkv, e.w16(p + "qkv_w"), e.dln_s, b_mn=True), 2.0 * M * C * 3 * Cl, ag_bytes),
        ("pull    AG copy of a [M,C] operand for a wgrad (K9 kernel)", lambda: e._pull(e._off_dx[0], e.full_a, side=False), 0.0, ag_bytes),
    ]
    rows = []
    for name, fn, flops, nbytes in cases:
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.reps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t.item()) * 1e3
        roof_us = max(flops / peak, nbytes / link) * 1e6
        rows.append({"op": name, "us": us, "mma_us": flops / peak * 1e6, "nvlink_us": nbytes / link * 1e6, "roofline_us": roof_us,
                     "fraction_of_roofline": roof_us / us, "gbytes_per_s": nbytes / (us * 1e-6) / 1e9})
    if rank == 0:
        print(f"K8 form: {mode}")
        print(f"TP x{W}, GPT-2 medium block, {args.batch} x 1024 tokens (M = {M}, shard {Ms} rows); peak {peak / 1e12:.0f} TFLOP/s, link 0.9 TB/s/dir")
        print(f"{'op':66s} {'us':>8s} {'MMA us':>8s} {'NVL us':>8s} {'roof':>6s} {'GB/s':>7s}")
        for r in rows:
            print(f"{r['op']:66s} {r['us']:8.1f} {r['mma_us']:8.1f} {r['nvlink_us']:8.1f} {r['fraction_of_roofline']:6.2f} {r['gbytes_per_s']:7.0f}")
        print(json.dumps({"world": W, "batch": args.batch, "rows": rows}))
    if args.step:
        del e
        torch.cuda.empty_cache()
        cfg = GPT2Config(**GPT2_PRESETS["gpt2-medium"])
        eng = GPT2TPEngine(cfg, batch_size=args.batch, seq_len=1024, device=dev, use_graph=True, security=True)
        g = torch.Generator().manual_seed(1)
        idx = torch.randint(0, cfg.vocab_size, (args.batch, 1024), generator=g).to(dev)
        eng.idx.copy_(idx); eng.tgt.copy_(idx)
        for _ in range(4):
            eng.run_step_device()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            eng.run_step_device()
        e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        if rank == 0:
            cb = eng.comm_bytes_per_step()
            print(json.dumps({"tp_engine_step": {"world": W, "global_batch": args.batch, "seq_len": 1024, "ms_per_step": ms,
                                                 "tokens_per_s": args.batch * 1024 / (ms / 1e3), "library_calls_per_step": eng.kernels_per_step(),
                                                 "nvlink_bytes_per_direction": cb, "nvlink_floor_ms": sum(cb.values()) / 0.9e12 * 1e3,
                                                 "loss": float(eng.loss_sum.item()) / eng.M * W}}))
    dist.barrier(); dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
