"""tcgen05 GEMM (K6) vs fp32 matmul of the same bf16 inputs: every operand layout and epilogue."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from trustworthy_dl.ops import reference_impl as R


def _mk(shape, dev, scale=1.0):
    return (torch.randn(*shape, device=dev) * scale).bfloat16()


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 256, 128), (384, 768, 256), (1000, 320, 192), (4096, 1024, 1024),
                                   (130, 96, 72)])
@pytest.mark.parametrize("block_n", [128, 256])
def test_gemm_nt_plain(M, N, K, block_n):
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(M + N + K)
    if K % 8:
        pytest.skip("K pitch must be 16-byte aligned")
    a, b = _mk((M, K), dev), _mk((N, K), dev)
    out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
    Kk.gemm(a, b, out, M, N, K, lda=K, ldb=K, ldc=N, block_n=block_n)
    ref = a.float() @ b.float().t()
    err = (out.float() - ref).abs().max().item()
    assert err <= 2e-2 * (ref.abs().max().item() + 1), f"max err {err}"


def test_gemm_dgrad_layout_b_mn_major():
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    M, N, Kin = 512, 768, 256                      # dy [M,N], w [N,Kin] -> dx [M,Kin]
    dy, w = _mk((M, N), dev), _mk((N, Kin), dev)
    dx = torch.zeros(M, Kin, device=dev, dtype=torch.bfloat16)
    Kk.linear_dgrad(dy, w, dx)
    ref = dy.float() @ w.float()
    assert (dx.float() - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)


@pytest.mark.parametrize("split_k", [1, 4])
def test_gemm_wgrad_layout_both_mn_major(split_k):
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(2)
    M, N, Kin = 2048, 384, 256                     # dy [M,N], x [M,Kin] -> dw [N,Kin] fp32 accumulate
    dy, x = _mk((M, N), dev), _mk((M, Kin), dev)
    dw = torch.ones(N, Kin, device=dev)
    Kk.linear_wgrad(dy, x, dw, split_k=split_k)
    ref = dy.float().t() @ x.float() + 1.0
    assert (dw - ref).abs().max().item() <= 1e-3 * (ref.abs().max().item() + 1)


def test_gemm_epilogues():
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    M, N, K = 256, 512, 128
    x, w = _mk((M, K), dev), _mk((N, K), dev, 0.2)
    bias = torch.randn(N, device=dev)
    res = _mk((M, N), dev)
    base = x.float() @ w.float().t() + bias
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, w, out, bias=bias)
    assert torch.allclose(out.float(), base, atol=5e-2, rtol=2e-2)
    Kk.linear_fwd(x, w, out, bias=bias.bfloat16())
    assert torch.allclose(out.float(), x.float() @ w.float().t() + bias.bfloat16().float(), atol=5e-2, rtol=2e-2)
    Kk.linear_fwd(x, w, out, bias=bias, epilogue=Kk.EPI_RESIDUAL, aux_in=res)
    assert torch.allclose(out.float(), base + res.float(), atol=6e-2, rtol=2e-2)
    pre = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, w, out, bias=bias, epilogue=Kk.EPI_GELU, aux_out=pre)
    assert torch.allclose(pre.float(), base, atol=5e-2, rtol=2e-2)
    assert torch.allclose(out.float(), R.gelu_tanh(pre), atol=3e-2, rtol=2e-2)
    # dgrad through GELU
    dy = _mk((M, K), dev)
    w2 = _mk((K, N), dev, 0.2)                     # Linear(N -> K): weight [K, N]
    dpre = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    Kk.linear_dgrad(dy, w2, dpre, epilogue=Kk.EPI_DGELU, aux_in=pre)
    ref = R.gelu_tanh_bwd(pre, dy.float() @ w2.float())
    assert torch.allclose(dpre.float(), ref, atol=5e-2, rtol=3e-2)


def test_gemm_lm_head_shapes():
    """N = 50304 (not a multiple of 256) and K = 50304 in the dgrad."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(4)
    M, C, Vp = 256, 256, 50304
    x, wte = _mk((M, C), dev), _mk((Vp, C), dev, 0.05)
    logits = torch.empty(M, Vp, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, wte, logits)
    ref = x.float() @ wte.float().t()
    assert (logits.float() - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)
    d = _mk((M, Vp), dev, 0.01)
    dx = torch.empty(M, C, device=dev, dtype=torch.bfloat16)
    Kk.linear_dgrad(d, wte, dx)
    ref = d.float() @ wte.float()
    assert (dx.float() - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)
    dw = torch.zeros(Vp, C, device=dev)
    Kk.linear_wgrad(d, x, dw)
    ref = d.float().t() @ x.float()
    assert (dw - ref).abs().max().item() <= 1e-3 * (ref.abs().max().item() + 1)


@pytest.mark.parametrize("W", [2, 4, 8])
def test_allgather_gemm_and_gemm_scatter_emulated_ranks(W):
    """K8 / K7 with all "ranks" on one GPU: shard pointers are local, the kernels are the real ones."""
    from trustworthy_dl.ops import nn as Kk, tp_ops
    dev = torch.device("cuda", 0)
    torch.manual_seed(W)
    shard_rows, Kd, N = 256, 256, 512
    M = shard_rows * W
    shards = [_mk((shard_rows, Kd), dev) for _ in range(W)]
    w = _mk((N, Kd), dev, 0.2)
    bias = torch.randn(N, device=dev)
    ref = torch.cat(shards).float() @ w.float().t() + bias
    for rank in (0, W - 1):
        out = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
        tp_ops.allgather_gemm([s.data_ptr() for s in shards], rank, shard_rows, w, out, bias=bias)
        assert torch.allclose(out.float(), ref, atol=6e-2, rtol=2e-2), (out.float() - ref).abs().max().item()
    # row-parallel: every rank holds a K-slice; partial products are scattered into the owners' slots
    Kfull = 128 * W
    x = _mk((M, Kfull), dev)
    w2 = _mk((N, Kfull), dev, 0.2)
    staging = [torch.zeros(W, shard_rows, N, device=dev, dtype=torch.bfloat16) for _ in range(W)]
    for rank in range(W):
        xs = x[:, rank * 128:(rank + 1) * 128].contiguous()
        ws = w2[:, rank * 128:(rank + 1) * 128].contiguous()
        tp_ops.gemm_scatter(xs, ws, [s.data_ptr() for s in staging], rank, shard_rows)
    ref2 = x.float() @ w2.float().t()
    res = _mk((shard_rows, N), dev)
    for owner in range(W):
        out = torch.empty(shard_rows, N, device=dev, dtype=torch.bfloat16)
        tp_ops.reduce_slots(staging[owner], W, out, residual=res)
        exp = ref2[owner * shard_rows:(owner + 1) * shard_rows] + res.float()
        assert torch.allclose(out.float(), exp, atol=0.15, rtol=3e-2), (out.float() - exp).abs().max().item()


@pytest.mark.parametrize("M,N,K,dgelu", [(1024, 512, 256, True), (2048, 768, 128, False), (1000, 320, 192, False),
                                         (256, 256, 128, False)])
def test_gemm_fused_column_sums(M, N, K, dgelu):
    """colsum += column sums of the bf16 output, from the staged epilogue panels (2-CTA path) or the stand-alone
    reduction (small M -> 1-CTA path); rows beyond M must not contribute."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(M + N)
    dy, w = _mk((M, K), dev), _mk((K, N), dev, 0.2)
    pre = _mk((M, N), dev) if dgelu else None
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    cs = torch.full((N,), 0.25, device=dev)
    Kk.linear_dgrad(dy, w, out, epilogue=Kk.EPI_DGELU if dgelu else Kk.EPI_NONE, aux_in=pre, colsum=cs)
    ref = dy.float() @ w.float()
    if dgelu:
        ref = R.gelu_tanh_bwd(pre, ref)
    assert torch.allclose(out.float(), ref, atol=5e-2, rtol=3e-2)
    want = out.float().sum(0) + 0.25
    assert torch.allclose(cs, want, atol=2e-3 * (M ** 0.5), rtol=1e-4), (cs - want).abs().max().item()


@pytest.mark.parametrize("M", [256, 1024])
def test_gemm_gelu_with_saved_derivative(M):
    """Epilogue 4 stores gelu'(pre) next to gelu(pre); epilogue 5 multiplies by it: together = the GELU backward."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(5 + M)
    N, K = 512, 128
    x, w = _mk((M, K), dev), _mk((N, K), dev, 0.2)
    bias = torch.randn(N, device=dev)
    pre = x.float() @ w.float().t() + bias
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    dg = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, w, out, bias=bias, epilogue=Kk.EPI_GELU_DG, aux_out=dg)
    assert torch.allclose(out.float(), R.gelu_tanh(pre), atol=3e-2, rtol=2e-2)
    dg_ref = R.gelu_tanh_bwd(pre, torch.ones_like(pre))
    assert torch.allclose(dg.float(), dg_ref, atol=2e-2, rtol=2e-2), (dg.float() - dg_ref).abs().max().item()
    dy = _mk((M, K), dev)
    w2 = _mk((K, N), dev, 0.2)
    dpre = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    cs = torch.zeros(N, device=dev)
    Kk.linear_dgrad(dy, w2, dpre, epilogue=Kk.EPI_MUL, aux_in=dg, colsum=cs)
    ref = (dy.float() @ w2.float()) * dg.float()
    assert torch.allclose(dpre.float(), ref, atol=5e-2, rtol=3e-2)
    assert torch.allclose(cs, dpre.float().sum(0), atol=2e-3 * (M ** 0.5), rtol=1e-4)


def test_gemm_lm_head_two_cta_grouped_rasterisation():
    """N = 50304 with M >= 512: the 2-CTA kernel with grouped tile order (197 n-blocks, the LM-head fast path)."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(11)
    M, C, Vp = 4096, 256, 50304                    # 16 m-pairs x 197 n-blocks: one full rasterisation group and change
    x, wte = _mk((M, C), dev), _mk((Vp, C), dev, 0.05)
    logits = torch.empty(M, Vp, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, wte, logits)
    ref = x.float() @ wte.float().t()
    assert (logits.float() - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)
    d = _mk((M, Vp), dev, 0.01)
    dx = torch.empty(M, C, device=dev, dtype=torch.bfloat16)
    Kk.linear_dgrad(d, wte, dx)                    # K = 50304 (786 k-blocks) through the 2-CTA kernel
    ref = d.float() @ wte.float()
    assert (dx.float() - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)


@pytest.mark.parametrize("N,Kin", [(1024, 4096), (4096, 1024), (3072, 1024), (1024, 1024)])
def test_gemm_wgrad_auto_split_k_on_gpt2_medium_shapes(N, Kin):
    """split_k = 0: the library picks the split that fills whole waves (what the engine always passes)."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(N + Kin)
    M = 8192                                       # tokens (the reduction dimension): 128 k-blocks
    dy, x = _mk((M, N), dev, 0.5), _mk((M, Kin), dev, 0.5)
    dw = torch.full((N, Kin), 0.5, device=dev)
    Kk.linear_wgrad(dy, x, dw, split_k=0)
    ref = dy.float().t() @ x.float() + 0.5
    assert (dw - ref).abs().max().item() <= 2e-3 * (ref.abs().max().item() + 1)


def test_gemm_full_lm_head_bench_shape_sampled_rows():
    """The shape bench.py runs (32768 x 50304 x 1024), checked on 64 random rows against fp32."""
    from trustworthy_dl.ops import nn as Kk
    dev = torch.device("cuda", 0)
    torch.manual_seed(12)
    M, C, Vp = 32768, 1024, 50304
    x, wte = _mk((M, C), dev), _mk((Vp, C), dev, 0.03)
    logits = torch.empty(M, Vp, device=dev, dtype=torch.bfloat16)
    Kk.linear_fwd(x, wte, logits)
    rows = torch.randint(0, M, (64,), device=dev)
    ref = x[rows].float() @ wte.float().t()
    got = logits[rows].float()
    assert (got - ref).abs().max().item() <= 2e-2 * (ref.abs().max().item() + 1)
    del logits
    torch.cuda.empty_cache()
