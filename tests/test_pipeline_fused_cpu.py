"""The fused pipeline strategy's schedule / transport / hand-off / reassignment code (parallel/pipeline_fused.py::train_step,
_ensure_groups, reassign) executed for real on CPU ranks over gloo, with a small stand-in for the stage engine.

The stand-in has the stage engine's buffer protocol (activation slots ``res[slot, 0..L]``, ``dy_in[parity]``, ``dx_out[parity]``,
``idx`` / ``tgt`` slots, ``_forward(slot)`` / ``_backward(slot, parity)``) and re-reads its input slot in the backward, like the
real engine re-reads the saved residual stream, so a receive posted too early or a message routed to the wrong stage or
micro-batch changes the numbers.  The oracle is the same chain of layers trained in one process by autograd; it does not
know about stages, ranks, slots or micro-batches.
"""
import pytest
import torch

from mp_utils import run_distributed

C, V, T, LAYERS, LR = 16, 32, 8, 8, 0.05


def _weights():
    g = torch.Generator().manual_seed(7)
    E = torch.randn(V, C, generator=g) * 0.5
    Ws = [torch.randn(C, C, generator=g) * (1.0 / C ** 0.5) for _ in range(LAYERS)]
    H = torch.randn(C, V, generator=g) * 0.3
    return E, Ws, H


def _batches(steps, batch):
    g = torch.Generator().manual_seed(11)
    return [{"input": torch.randint(0, V, (batch, T), generator=g), "target": torch.randint(0, V, (batch, T), generator=g)}
            for _ in range(steps)]


def _oracle(steps, batch, skip=None):
    """``skip``: {step index: (a, b)} -> blocks [a, b) keep their weights in that step (a dropped stage update)."""
    E, Ws, H = _weights()
    params = [E] + Ws + [H]
    for p in params:
        p.requires_grad_(True)
    losses = []
    for b in _batches(steps, batch):
        h = E[b["input"]]
        for W in Ws:
            h = torch.tanh(h @ W)
        loss = torch.nn.functional.cross_entropy((h @ H).view(-1, V), b["target"].view(-1))
        loss.backward()
        frozen = set(range(*skip[len(losses)])) if skip and len(losses) in skip else set()
        with torch.no_grad():
            for j, p in enumerate(params):
                if j - 1 not in frozen or j == 0 or j == len(params) - 1:
                    p -= LR * p.grad
                p.grad = None
        losses.append(float(loss.detach()))
    val = 0.0
    with torch.no_grad():
        for b in _batches(2, batch):
            h = E[b["input"]]
            for W in Ws:
                h = torch.tanh(h @ W)
            val += float(torch.nn.functional.cross_entropy((h @ H).view(-1, V), b["target"].view(-1))) / 2
    return losses, float(sum(p.detach().double().sum() for p in params)), val


class _Reducer:
    def __init__(self, ranks):
        self.ranks = list(ranks)
        self.trust = torch.ones(len(ranks))
        self.status = torch.zeros(len(ranks), dtype=torch.int32)

    def set_group(self, ranks):
        self.ranks = list(ranks)


class _StageEngine:
    """Layers [a, b) of the chain (+ embedding / head) behind the stage engine's buffer protocol."""

    def __init__(self, sid, a, b, has_embed, has_head, slots, mb, micro, dp_ranks):
        import torch.distributed as dist
        self.sid, self.L, self.slots, self.mb, self.micro = sid, b - a, slots, mb, micro
        self.has_embed, self.has_head = has_embed, has_head
        E, Ws, H = _weights()
        self.params = ([E] if has_embed else []) + Ws[a:b] + ([H] if has_head else [])
        for p in self.params:
            p.requires_grad_(True)
        self.res = torch.zeros(slots, self.L + 1, mb * T, C)
        self.dy_in = torch.zeros(2, mb * T, C)
        self.dx_out = torch.zeros(2, mb * T, C)
        self.idx = torch.zeros(slots, mb, T, dtype=torch.long)
        self.tgt = torch.zeros(slots, mb, T, dtype=torch.long)
        self.attack_scale = torch.ones(1)
        self.loss_sum = torch.zeros(1)
        self.grad_sumsq = torch.zeros(1, dtype=torch.float64)
        self.M = mb * T
        self.rank = dist.get_rank()
        self.reducer = _Reducer(dp_ranks)
        self.world = len(dp_ranks)

    def begin_step(self):
        self.loss_sum.zero_()

    def _chain(self, slot):
        if self.has_embed:
            x, h = None, self.params[0][self.idx[slot].view(-1)]
        else:
            x = self.res[slot, 0].clone().requires_grad_(True)
            h = x
        ws = self.params[1 if self.has_embed else 0:len(self.params) - (1 if self.has_head else 0)]
        for W in ws:
            h = torch.tanh(h @ W)
        return x, h

    def _forward(self, slot):
        with torch.no_grad():
            _, h = self._chain(slot)
            self.res[slot, self.L].copy_(h)

    def _backward(self, slot, k):
        x, h = self._chain(slot)                       # re-reads the input slot, like the saved residual stream
        if self.has_head:
            ce = torch.nn.functional.cross_entropy(h @ self.params[-1], self.tgt[slot].view(-1), reduction="sum")
            self.loss_sum += ce.detach()
            (ce / (self.M * self.micro)).backward()
        else:
            h.backward(self.dy_in[k])
        if x is not None:
            self.dx_out[k].copy_(x.grad)

    def head_loss_sum(self, slot, targets):
        with torch.no_grad():
            return torch.nn.functional.cross_entropy(self.res[slot, self.L] @ self.params[-1], targets.reshape(-1),
                                                     reduction="sum").double()

    device = torch.device("cpu")

    def state_dict(self):
        return {"flat": torch.cat([p.detach().reshape(-1) for p in self.params])}

    def load_state_dict(self, sd):
        off = 0
        with torch.no_grad():
            for p in self.params:
                p.copy_(sd["flat"][off:off + p.numel()].view_as(p))
                off += p.numel()

    def export_to_module(self, model):
        pass

    def flat_grad(self):
        return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in self.params])

    def reduce_and_update(self):
        import torch.distributed as dist
        g = self.flat_grad()
        self.grad_sumsq[0] = float(g.double().pow(2).sum())
        if self.world > 1:                              # mean over the replicas of this stage (tagged by stage id)
            peers = [r for r in self.reducer.ranks if r != self.rank]
            got = [torch.empty_like(g) for _ in peers]
            reqs = [dist.isend(g.clone(), p, tag=100 + self.sid) for p in peers] + \
                   [dist.irecv(b, p, tag=100 + self.sid) for b, p in zip(got, peers)]
            for r in reqs:
                r.wait()
            g = (g + sum(got)) / self.world
        off = 0
        with torch.no_grad():
            for p in self.params:
                n = p.numel()
                p -= LR * g[off:off + n].view_as(p)
                p.grad = None
                off += n


def _make_strategy(trainer):
    import torch.distributed as dist
    from trustworthy_dl.ops.reference_impl import tensor_statistics
    from trustworthy_dl.parallel.pipeline_fused import FusedPipelineStrategy

    class Strategy(FusedPipelineStrategy):
        def _alloc_shared(self, cfg):
            pass

        def _tensor_stats(self, x, out):
            out.copy_(tensor_statistics(x))

        def _local_grad_sumsq(self, eng):
            eng.grad_sumsq[0] = float(eng.flat_grad().double().pow(2).sum())

        def _make_engine(self, sid, load, ranks=None):
            a, b = self.ranges[sid]
            return _StageEngine(sid, a, b, sid == 0, sid == self.S - 1, self._slots(sid), self.mb, self.micro,
                                ranks if ranks is not None else self._dp_ranks(sid))

        def _migrate(self, sid, src, dst):
            if self.rank == src:
                eng = self.engines.pop(sid)
                dist.send(torch.cat([p.detach().reshape(-1) for p in eng.params]), dst, tag=900 + sid)
            elif self.rank == dst:
                eng = self._make_engine(sid, False, [dst if r == src else r for r in self._dp_ranks(sid)])
                flat = torch.empty(sum(p.numel() for p in eng.params))
                dist.recv(flat, src, tag=900 + sid)
                off = 0
                with torch.no_grad():
                    for p in eng.params:
                        p.copy_(flat[off:off + p.numel()].view_as(p))
                        off += p.numel()
                self.engines[sid] = eng
            dist.barrier()

    return Strategy(trainer)


def _worker(rank, world, dp, virtual, micro, batch, steps, moves, attack_step=None):
    from trustworthy_dl import AdversarialAttacker, AttackConfig, DistributedTrainer, TrainingConfig
    att = None
    if attack_step is not None:
        att = AdversarialAttacker(AttackConfig(attack_types=["gradient_poisoning"], target_nodes=[1], intensity=0.5,
                                               start_step=attack_step))
        att.activate_attacks()
    cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=batch, num_nodes=world,
                         learning_rate=LR, checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=micro,
                         dp_size=dp, pp_virtual=virtual, seq_len=T, gradient_verification_enabled=attack_step is not None,
                         model_kwargs={"n_layer": LAYERS, "n_positions": T})
    t = DistributedTrainer(cfg, attacker=att)
    t._rank, t._world = rank, world
    t._backend_impl = _make_strategy(t)
    t._backend_impl.setup()
    st = t._backend_impl
    losses, flows = [], []
    for i, b in enumerate(_batches(steps, batch)):
        for at, src in moves:
            if at == i:
                t.global_step = i
                assert st.reassign(src) is not None
        losses.append(t.train_step(b))
        flows.append(len(st._pair_groups))
    val = st.validate(_batches(2, batch))
    wsum = float(sum(p.detach().double().sum() for e in st.engines.values() for p in e.params))
    return {"losses": losses, "val": val, "owner": list(t.stage_map.owner), "stages": sorted(st.engines), "wsum": wsum, "flows": flows,
            "ranges": list(st.ranges), "attacks": len(t.attack_history),
            "attack_log": [(a["node_id"], a["step"], a["attack_type"]) for a in t.attack_history],
            "reassign": [(r["from_node"], r["step"], r["stages"]) for r in t.reassignment_history]}


@pytest.mark.parametrize("world,dp,virtual,micro,batch,moves", [
    (2, 1, 2, 4, 8, [(3, 1)]),             # interleaved stages {0,2} / {1,3}; then rank 0 hosts everything (local hand-offs only)
    (3, 1, 2, 6, 12, [(2, 2)]),            # six interleaved stages on three ranks, the rank with the head gives its stages up
    (4, 2, 2, 4, 16, [(2, 3), (4, 0)]),    # interleaved PP x DP, one move in each replica
    (8, 2, 1, 4, 16, [(3, 2)]),            # BASELINE config 5 layout (PP x4 x DP x2) with a mid-run move
])
def test_fused_pipeline_transport_matches_single_process(world, dp, virtual, micro, batch, moves):
    steps = 6
    ref_losses, ref_wsum, ref_val = _oracle(steps, batch)
    res = run_distributed(_worker, world, dp, virtual, micro, batch, steps, moves, timeout=300)
    for r in res:
        assert r["losses"] == pytest.approx(ref_losses, rel=2e-4, abs=2e-5), (r["losses"], ref_losses)
        assert r["owner"] == res[0]["owner"] and r["attacks"] == 0
        assert r["val"] == pytest.approx(ref_val, rel=2e-4)          # forward-only pass over the same transport
    for at, src in moves:
        assert src not in res[0]["owner"]                          # the rank gave up everything it hosted
    # every stage of every replica is hosted exactly once, and the hosted weights add up to the oracle's (x dp replicas)
    hosted = sorted(s for r in res for s in r["stages"])
    S = len(res[0]["ranges"])
    assert hosted == sorted(list(range(S)) * dp)
    assert sum(r["wsum"] for r in res) == pytest.approx(dp * ref_wsum, rel=1e-5)
    assert res[0]["flows"][-1] >= res[0]["flows"][0]


def test_fused_pipeline_poisoned_stage_is_flagged_moved_and_its_update_dropped():
    """Rank 1 ships an 11x gradient at step 7: every rank flags it in that step, its stage moves (parameters intact), the
    poisoned update is dropped -- by the old owner AND by the new one, which adopts the stage without gradients -- and
    training continues on the new layout exactly like the single-process chain with that one update skipped."""
    from trustworthy_dl.parallel.pipeline_fused import split_layers
    steps, batch = 10, 8
    a, b = split_layers(LAYERS, 3)[1]
    ref_losses, ref_wsum, ref_val = _oracle(steps, batch, skip={6: (a, b)})
    res = run_distributed(_worker, 3, 1, 1, 4, batch, steps, [], 7, timeout=300)
    for r in res:
        assert r["attack_log"] == [(1, 7, "gradient_poisoning")]
        assert r["reassign"] == [(1, 7, [1])]
        assert 1 not in r["owner"]
        assert r["losses"] == pytest.approx(ref_losses, rel=2e-4, abs=2e-5), (r["losses"], ref_losses)
        assert r["val"] == pytest.approx(ref_val, rel=2e-4)
    assert sum(r["wsum"] for r in res) == pytest.approx(ref_wsum, rel=1e-5)


def _resume_worker(rank, world, ckpt_dir, phase):
    """'a': 4 steps with a move before step 2, checkpoint, 2 more steps.  'b': fresh ranks resume the step-4 files."""
    from trustworthy_dl import DistributedTrainer, TrainingConfig
    cfg = TrainingConfig(model_name="gpt2-tiny", dataset_name="tiny-tokens", batch_size=8, num_nodes=world, learning_rate=LR,
                         checkpoint_interval=0, device="cpu", parallelism="pipeline", micro_batches=4, dp_size=1, pp_virtual=2,
                         seq_len=T, gradient_verification_enabled=False, checkpoint_dir=ckpt_dir,
                         model_kwargs={"n_layer": LAYERS, "n_positions": T})
    t = DistributedTrainer(cfg)
    t._rank, t._world = rank, world
    t._backend_impl = _make_strategy(t)
    t._backend_impl.setup()
    st = t._backend_impl
    batches = _batches(6, 8)
    if phase == "a":
        for i, b in enumerate(batches[:4]):
            if i == 2:
                t.global_step = i
                assert st.reassign(1) is not None
            t.train_step(b)
        t.save_checkpoint()
    else:
        t.load_checkpoint()
    tail = [t.train_step(b) for b in batches[4:]]
    return {"tail": tail, "owner": list(t.stage_map.owner), "stages": sorted(st.engines), "step": t.global_step}


def test_fused_pipeline_resume_after_reassignment(tmp_path):
    ref_losses, _, _ = _oracle(6, 8)
    a = run_distributed(_resume_worker, 3, str(tmp_path), "a", timeout=300)
    b = run_distributed(_resume_worker, 3, str(tmp_path), "b", timeout=300)
    assert a[0]["owner"] == b[0]["owner"] and 1 not in b[0]["owner"] and b[1]["stages"] == []
    assert [r["stages"] for r in a] == [r["stages"] for r in b]
    for r in a + b:
        assert r["tail"] == pytest.approx(ref_losses[4:], rel=2e-4)
    assert b[0]["step"] == 6
