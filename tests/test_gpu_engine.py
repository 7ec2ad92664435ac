"""Fused GPT-2 engine vs the nn.Module definition (autograd) on a small config."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tiny():
    from trustworthy_dl.models.gpt2 import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(0)
    cfg = GPT2Config(vocab_size=1000, padded_vocab_size=1024, n_positions=128, n_embd=128, n_layer=2, n_head=2)
    return cfg, GPT2LMHeadModel(cfg)


def test_engine_logits_and_grads_match_module():
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    dev = torch.device("cuda", 0)
    cfg, model = _tiny()
    model = model.to(dev)
    eng = GPT2Engine.from_module(model, batch_size=2, seq_len=128, device=dev, use_graph=False, security=False)
    idx = torch.randint(0, 1000, (2, 128), device=dev)
    tgt = torch.randint(0, 1000, (2, 128), device=dev)
    logits = eng.forward_logits(idx)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ref = model(idx).float()
    assert torch.allclose(logits, ref, atol=8e-2, rtol=5e-2), (logits - ref).abs().max().item()
    # gradients: run fwd+bwd only
    eng.idx[0].copy_(idx); eng.tgt[0].copy_(tgt)
    eng.loss_sum.zero_(); eng.g32.zero_()
    eng._forward(); eng._backward()
    torch.cuda.synchronize()
    loss = model.loss(model(idx), tgt)
    loss.backward()
    assert eng.loss_sum.item() / eng.M == pytest.approx(loss.item(), rel=2e-2)
    ref_grads = {n: t.grad for n, t in eng._module_tensors(model).items()}
    assert set(ref_grads) == set(eng.offsets)
    for name in eng.offsets:                                   # every parameter tensor of the model
        g, r = eng.grad32(name), ref_grads[name].float()
        denom = r.abs().max().item() + 1e-6
        err = (g - r).abs().max().item() / denom
        assert err < 8e-2, f"{name}: rel err {err}"


def _small(n_layer=4):
    from trustworthy_dl.models.gpt2 import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(1)
    cfg = GPT2Config(vocab_size=2000, padded_vocab_size=2048, n_positions=256, n_embd=768, n_layer=n_layer, n_head=12)
    return cfg, GPT2LMHeadModel(cfg)


def test_engine_ten_steps_match_torch_adamw_at_gpt2_small_width():
    """Engine vs GPT2LMHeadModel + torch.optim.AdamW for 10 steps at GPT-2-small width: the loss curves stay within 1 %
    and every parameter tensor ends up where the stock optimizer put it (bf16-compute tolerance)."""
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    dev = torch.device("cuda", 0)
    cfg, model = _small()
    model = model.to(dev)
    B, T, lr = 4, 256, 1e-4
    eng = GPT2Engine.from_module(model, batch_size=B, seq_len=T, lr=lr, weight_decay=0.01, device=dev, use_graph=True, security=True)
    opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=0.01, betas=(0.9, 0.999), eps=1e-8)
    g = torch.Generator().manual_seed(5)
    data = [(torch.randint(0, 2000, (B, T), generator=g).to(dev), torch.randint(0, 2000, (B, T), generator=g).to(dev)) for _ in range(3)]
    ours, ref = [], []
    for i in range(10):
        idx, tgt = data[i % 3]
        ours.append(eng.train_step(idx, tgt))
        with torch.autocast("cuda", dtype=torch.bfloat16):
            logits = model(idx)
        loss = model.loss(logits, tgt)
        loss.backward()
        opt.step(); opt.zero_grad(set_to_none=True)
        ref.append(float(loss.item()))
    assert ours == pytest.approx(ref, rel=1e-2), (ours, ref)
    assert ours[-1] < ours[0]
    want = eng._module_tensors(model)
    for name in eng.offsets:
        a, b = eng.w32(name), want[name].detach().float()
        # 10 Adam steps move every weight by <= 10*lr; where a gradient is rounding-sized Adam's step is +-lr with either
        # sign, so single elements may end up to 2*10*lr apart while the bulk agrees closely
        d = (a - b).abs()
        assert d.max().item() <= 20 * lr and d.mean().item() < 2 * lr, (name, d.max().item(), d.mean().item())


def test_overlapped_and_serial_schedules_agree():
    """Aggregation + AdamW issued per segment on the side stream under backward == the same work issued afterwards."""
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    dev = torch.device("cuda", 0)
    cfg, _ = _tiny()
    g = torch.Generator().manual_seed(0)
    idx = torch.randint(0, 1000, (2, 128), generator=g).to(dev)
    tgt = torch.randint(0, 1000, (2, 128), generator=g).to(dev)
    out = []
    for overlap, graph in ((False, False), (True, False), (True, True)):
        eng = GPT2Engine(cfg, batch_size=2, seq_len=128, lr=1e-3, device=dev, use_graph=graph, security=True, overlap=overlap,
                         n_buckets=2)
        losses = [eng.train_step(idx, tgt) for _ in range(6)]
        snap = eng.security_snapshot()
        out.append((losses, eng.p32.clone(), snap["stats"][0, 0].item(), snap["trust"].item()))
        assert len(eng.segments) == 3
    # split-K wgrads accumulate with fp32 atomics, so two runs agree to rounding, not bit for bit
    for o in out[1:]:
        assert o[0] == pytest.approx(out[0][0], rel=2e-3)
        # (Adam turns a rounding-sized difference in a near-zero gradient into a step of up to lr, so compare the bulk)
        d = (o[1] - out[0][1]).abs()
        assert d.max().item() < 3e-3 and d.mean().item() < 2e-5
        assert o[2] == pytest.approx(out[0][2], rel=1e-2) and o[3] == pytest.approx(out[0][3], abs=1e-4)


def test_stage_engines_chain_to_the_full_model():
    """Two stage engines (embedding + blocks 0-1 | blocks 2-3 + ln_f + untied head) fed through their p2p buffers give the
    loss and the gradients of one engine that holds the whole (untied) model, micro-batch accumulation included."""
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    dev = torch.device("cuda", 0)
    cfg, model = _small()
    model.untie_head()
    model = model.to(dev)
    B, T, MB = 2, 256, 2
    kw = dict(batch_size=B, seq_len=T, device=dev, use_graph=False, security=False)
    full = GPT2Engine.from_module(model, tied_head=False, micro_slots=MB, micro_batches=MB, **kw)
    s0 = GPT2Engine.from_module(model, layers=(0, 2), has_head=False, micro_slots=MB, micro_batches=MB, **kw)
    s1 = GPT2Engine.from_module(model, layers=(2, 4), has_embed=False, tied_head=False, micro_slots=MB, micro_batches=MB, **kw)
    g = torch.Generator().manual_seed(9)
    idx = torch.randint(0, 2000, (MB, B, T), generator=g).to(dev)
    tgt = torch.randint(0, 2000, (MB, B, T), generator=g).to(dev)
    for e in (full, s0, s1):
        e.begin_step()
    full.idx.copy_(idx); full.tgt.copy_(tgt); s0.idx.copy_(idx); s1.tgt.copy_(tgt)
    for m in range(MB):
        full._forward(m)
        s0._forward(m)
        s1.res[m, 0].copy_(s0.res[m, s0.L])                    # what NCCL p2p does between the two ranks
        s1._forward(m)
    for m in range(MB):
        full._backward(m)
        s1._backward(m, m % 2)
        s0.dy_in[m % 2].copy_(s1.dx_out[m % 2])
        s0._backward(m, m % 2)
    torch.cuda.synchronize()
    assert s1.loss_sum.item() == pytest.approx(full.loss_sum.item(), rel=1e-3)
    for eng in (s0, s1):
        for name in eng.offsets:
            a, b = eng.grad32(name), full.grad32(name)
            denom = b.abs().max().item() + 1e-9
            assert (a - b).abs().max().item() / denom < 2e-2, name


@pytest.mark.parametrize("use_graph", [False, True])
def test_engine_trains_and_security_state(use_graph):
    from trustworthy_dl.engine.gpt2_engine import GPT2Engine
    dev = torch.device("cuda", 0)
    cfg, _ = _tiny()
    eng = GPT2Engine(cfg, batch_size=2, seq_len=128, lr=3e-3, device=dev, use_graph=use_graph, security=True)
    g = torch.Generator().manual_seed(0)
    idx = torch.randint(0, 1000, (2, 128), generator=g).pin_memory()
    tgt = torch.randint(0, 1000, (2, 128), generator=g).pin_memory()
    losses = [eng.train_step(idx, tgt) for _ in range(12)]
    assert losses[-1] < losses[0] - 1.0, losses
    snap = eng.security_snapshot()
    assert snap["status"].tolist() == [0] and 0.85 < snap["trust"].item() <= 1.0
    assert snap["stats"][0, 0].item() > 0 and snap["act_stats"][1].item() > 0
    assert eng.kernels_per_step() > 50


def test_smoke_entry():
    import __graft_entry__ as ge
    ge.smoke()
