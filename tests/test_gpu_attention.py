"""tcgen05 causal attention (K12) forward/backward vs the fp32 PyTorch oracle."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from trustworthy_dl.ops import reference_impl as R


def _split(qkv, B, T, H):
    C = H * 64
    q, k, v = qkv.float().view(B, T, 3, H, 64).unbind(2)
    return q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 256, 2), (1, 1024, 4), (3, 384, 3)])
def test_attention_forward(B, T, H):
    from trustworthy_dl.ops import nn as K
    dev = torch.device("cuda", 0)
    torch.manual_seed(B * 1000 + T + H)
    C = H * 64
    qkv = (torch.randn(B * T, 3 * C, device=dev)).bfloat16()
    out = torch.empty(B * T, C, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(B * H * T, device=dev)
    K.attention_fwd(qkv, out, lse, B, T, H)
    q, k, v = _split(qkv, B, T, H)
    o_ref, lse_ref = R.causal_attention_fwd(q, k, v)
    got = out.float().view(B, T, H, 64).transpose(1, 2)
    assert torch.allclose(got, o_ref, atol=3e-2, rtol=3e-2), (got - o_ref).abs().max().item()
    assert torch.allclose(lse.view(B, H, T), lse_ref, atol=2e-2, rtol=1e-3)


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 256, 2), (1, 512, 2), (2, 1024, 2), (1, 2048, 1)])
def test_attention_backward(B, T, H):
    from trustworthy_dl.ops import nn as K
    dev = torch.device("cuda", 0)
    torch.manual_seed(7 + T)
    C = H * 64
    qkv = torch.randn(B * T, 3 * C, device=dev).bfloat16()
    out = torch.empty(B * T, C, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(B * H * T, device=dev)
    K.attention_fwd(qkv, out, lse, B, T, H)
    d_out = torch.randn(B * T, C, device=dev).bfloat16()
    dqkv = torch.zeros(B * T, 3 * C, device=dev, dtype=torch.bfloat16)
    delta = torch.empty(B * H * T, device=dev)
    K.attention_bwd(qkv, out, d_out, lse, delta, dqkv, B, T, H)
    q, k, v = _split(qkv, B, T, H)
    o_ref, lse_ref = R.causal_attention_fwd(q, k, v)
    do = d_out.float().view(B, T, H, 64).transpose(1, 2)
    dq, dk, dv = R.causal_attention_bwd(q, k, v, o_ref, lse_ref, do)
    gq, gk, gv = _split(dqkv, B, T, H)
    for name, g, r in (("dq", gq, dq), ("dk", gk, dk), ("dv", gv, dv)):
        err = (g - r).abs().max().item()
        assert err <= 4e-2 * (r.abs().max().item() + 1), f"{name} max err {err}"


def test_attention_backward_fused_bias_grad():
    """The qkv bias gradient accumulated from the accumulator write-out equals the column sums of dqkv."""
    from trustworthy_dl.ops import nn as K
    dev = torch.device("cuda", 0)
    torch.manual_seed(21)
    B, T, H = 2, 384, 3
    C = H * 64
    qkv = torch.randn(B * T, 3 * C, device=dev).bfloat16()
    out = torch.empty(B * T, C, device=dev, dtype=torch.bfloat16)
    lse = torch.empty(B * H * T, device=dev)
    K.attention_fwd(qkv, out, lse, B, T, H)
    d_out = torch.randn(B * T, C, device=dev).bfloat16()
    dqkv = torch.zeros(B * T, 3 * C, device=dev, dtype=torch.bfloat16)
    delta = torch.empty(B * H * T, device=dev)
    dbias = torch.full((3 * C,), 0.5, device=dev)             # accumulates on top of what is there
    K.attention_bwd(qkv, out, d_out, lse, delta, dqkv, B, T, H, dbias=dbias)
    ref = dqkv.float().sum(0) + 0.5
    assert torch.allclose(dbias, ref, atol=2e-3, rtol=1e-4), (dbias - ref).abs().max().item()
    dqkv2 = torch.zeros_like(dqkv)
    K.attention_bwd(qkv, out, d_out, lse, delta, dqkv2, B, T, H)          # no bias pointer: same dqkv
    assert torch.equal(dqkv, dqkv2)
