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
    eng.idx.copy_(idx); eng.tgt.copy_(tgt)
    eng.loss_sum.zero_(); eng.g32.zero_()
    eng._forward(); eng._backward()
    torch.cuda.synchronize()
    loss = model.loss(model(idx), tgt)
    loss.backward()
    assert eng.loss_sum.item() / eng.M == pytest.approx(loss.item(), rel=2e-2)
    ref_grads = {n: t.grad for n, t in eng._module_tensors(model).items()}
    for name in ("h1.proj_w", "h1.fc_w", "h0.qkv_w", "h0.o_w", "h0.ln1_w", "h1.fc_b", "h0.qkv_b", "lnf_w", "wpe", "wte"):
        g, r = eng.grad32(name), ref_grads[name].float()
        denom = r.abs().max().item() + 1e-6
        err = (g - r).abs().max().item() / denom
        assert err < 8e-2, f"{name}: rel err {err}"


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
