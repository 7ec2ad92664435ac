"""Multi-GPU checks (gpurun --gpus N): symmetric-memory trust-weighted all-reduce vs the oracle, and
data-parallel engine replicas staying in lock-step while one rank is poisoned."""
import os
import socket
import traceback

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    try:
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        ret[rank] = ("ok", fn(rank, world))
    except Exception:
        ret[rank] = ("err", traceback.format_exc())
    finally:
        try:
            dist.destroy_process_group()
        except Exception:
            pass


def _run(fn, world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_entry, args=(r, world, port, fn, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
    for p in procs:
        if p.is_alive():
            p.kill()
            pytest.fail("multi-GPU worker timed out")
    out = []
    for r in range(world):
        st, val = ret.get(r, ("err", "no result"))
        assert st == "ok", f"rank {r}:\n{val}"
        out.append(val)
    return out


def _reducer_worker(rank, world):
    import torch.distributed as dist
    from trustworthy_dl.ops import reference_impl as R
    from trustworthy_dl.parallel.symm import SymmetricTrustReducer
    dev = torch.device("cuda", rank)
    N = 8 * world * 4096
    red = SymmetricTrustReducer(N, torch.bfloat16, dev)
    results = []
    for step in range(3):
        g = torch.Generator().manual_seed(100 + step)
        base = torch.randn(N, generator=g)
        mine = None
        alls = []
        for r in range(world):
            v = (base + 0.3 * torch.randn(N, generator=g)).bfloat16()
            if step == 2 and r == world - 1:
                v = (v.float() * 10).bfloat16()                      # poisoned rank in the last step
            alls.append(v)
        red.buffer.copy_(alls[rank].to(dev))
        red.enqueue_local_norm()
        red.enqueue_reduce(update_trust=True)
        torch.cuda.synchronize()
        norms = torch.stack([a.float().norm() for a in alls])
        tw = torch.full((world,), 1.0)
        coeff = R.robust_coefficients(norms, tw / world)
        ref, stats = R.trust_weighted_reduce([a.float() for a in alls], coeff, 0.0, norms)
        err = (red.buffer.float().cpu() - ref).abs().max().item()
        results.append({"err": err, "coeff": red.coeff.cpu().tolist(), "ref_coeff": coeff.tolist(),
                        "cos": red.stats[:, 3].cpu().tolist(), "ref_cos": stats[:, 3].tolist(),
                        "flags": red.flags.cpu().tolist(), "trust": red.trust.cpu().tolist()})
    dist.barrier()
    return results


def test_symmetric_trust_allreduce_matches_oracle():
    world = torch.cuda.device_count()
    res = _run(_reducer_worker, world)
    for rank_res in res:
        for step, r in enumerate(rank_res):
            assert r["err"] < 0.13, r                       # <= 2 bf16 ulps at |x| ~ 8
            assert r["coeff"] == pytest.approx(r["ref_coeff"], abs=2e-3)
            assert r["cos"] == pytest.approx(r["ref_cos"], abs=2e-2)
        last = rank_res[2]
        assert last["coeff"][-1] == 0.0 and last["flags"][-1] != 0          # the x10 rank is excluded in the same step
        assert last["trust"][-1] < 0.2 < last["trust"][0]
    assert all(r[2]["trust"] == pytest.approx(res[0][2]["trust"]) for r in res)   # replicated decisions


def _engine_worker(rank, world):
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    from trustworthy_dl.models.gpt2 import GPT2Config
    dev = torch.device("cuda", rank)
    cfg = GPT2Config(vocab_size=1000, padded_vocab_size=1024, n_positions=128, n_embd=128, n_layer=2, n_head=2)
    eng = GPT2Engine(cfg, batch_size=2, seq_len=128, lr=1e-3, device=dev, use_graph=True, security=True)
    g = torch.Generator().manual_seed(rank)
    idx = torch.randint(0, 1000, (2, 128), generator=g)
    losses = [eng.train_step(idx, idx.roll(-1, 1)) for _ in range(8)]
    torch.cuda.synchronize()
    snap = eng.security_snapshot()
    return {"losses": losses, "psum": float(eng.p32.double().sum().item()), "trust": snap["trust"].tolist(),
            "coeff": snap["coeff"].tolist()}


def test_data_parallel_engine_replicas_stay_in_sync():
    world = torch.cuda.device_count()
    res = _run(_engine_worker, world)
    assert len({round(r["psum"], 4) for r in res}) == 1                     # identical parameters on every rank
    assert all(r["losses"][-1] < r["losses"][0] for r in res)
    assert res[0]["coeff"] == pytest.approx([1.0 / world] * world, abs=0.05)
