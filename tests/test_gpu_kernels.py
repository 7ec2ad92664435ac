"""sm_100a kernels vs plain-PyTorch fp32 oracles (run on the B200 box: pytest -m gpu)."""
import ctypes
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from trustworthy_dl.ops import reference_impl as R


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


@pytest.mark.parametrize("dtype,n", [(torch.bfloat16, 1 << 20), (torch.float32, 300_001), (torch.bfloat16, 4097),
                                     (torch.float16, 65_536), (torch.bfloat16, 8 * 1024 * 1024)])
def test_tensor_statistics(dev, dtype, n):
    from trustworthy_dl.ops.stats import tensor_statistics
    g = torch.Generator(device="cpu").manual_seed(n)
    x = (torch.randn(n, generator=g) * 1.7 + 0.4).to(dtype).to(dev)
    got = tensor_statistics(x).cpu()
    ref = R.tensor_statistics(x.cpu().float())
    names = R.STAT_NAMES
    for i, name in enumerate(names):
        tol = 2e-4 if name in ("skewness", "kurtosis") else 1e-5
        assert got[i].item() == pytest.approx(ref[i].item(), rel=tol, abs=tol), name
    # exact order statistics
    for i in (2, 3, 4, 7, 8, 11):
        assert got[i].item() == pytest.approx(ref[i].item(), rel=1e-6, abs=1e-7), names[i]


def test_statistics_large_offset_and_detector_decisions(dev):
    """|mean| >> std conditioning + the detector makes the same decisions from device stats."""
    from trustworthy_dl import AttackDetector
    from trustworthy_dl.ops.stats import tensor_statistics
    x = (torch.randn(1 << 18) * 0.01 + 100.0).to(dev)
    got, ref = tensor_statistics(x).cpu(), R.tensor_statistics(x.cpu())
    assert got[1].item() == pytest.approx(ref[1].item(), rel=1e-3)
    torch.manual_seed(0)
    a, b = AttackDetector(), AttackDetector(stats_fn=R.tensor_statistics)
    for step in range(30):
        t = torch.randn(4, 64, 128)
        assert a.detect_output_anomaly(t.to(dev), 0, step) == b.detect_output_anomaly(t, 0, step) is False
    t = torch.randn(4, 64, 128) * 3
    assert a.detect_output_anomaly(t.to(dev), 0, 30) == b.detect_output_anomaly(t, 0, 30) is True


@pytest.mark.parametrize("M,C", [(64, 1024), (33, 256), (16, 768), (8, 1280)])
def test_layernorm_fwd_bwd(dev, M, C):
    from trustworthy_dl.ops import nn as K
    torch.manual_seed(0)
    x = torch.randn(M, C, device=dev).bfloat16()
    w = torch.randn(C, device=dev) * 0.5 + 1
    b = torch.randn(C, device=dev) * 0.1
    y = torch.empty_like(x); mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
    K.layernorm_fwd(x, w, b, y, mean, rstd)
    yr, mr, rr = R.layernorm_fwd(x, w, b)
    assert torch.allclose(y.float(), yr.float(), atol=2e-2, rtol=2e-2)
    assert torch.allclose(mean, mr, atol=1e-4) and torch.allclose(rstd, rr, rtol=1e-3)
    dy = torch.randn(M, C, device=dev).bfloat16()
    dres = torch.randn(M, C, device=dev).bfloat16()
    dx = torch.empty_like(x); dw = torch.zeros(C, device=dev); db = torch.zeros(C, device=dev)
    K.layernorm_bwd(dy, x, w, mean, rstd, dres, dx, dw, db)
    dxr, dwr, dbr = R.layernorm_bwd(dy, x, w, mr, rr)
    assert torch.allclose(dx.float(), dxr + dres.float(), atol=3e-2, rtol=3e-2)
    assert torch.allclose(dw, dwr, atol=1e-2 * math.sqrt(M), rtol=1e-2)
    assert torch.allclose(db, dbr, atol=1e-2 * math.sqrt(M), rtol=1e-2)
    # fused column sums of the written dx (bias gradient of the layer below); accumulates; no dres
    dx2 = torch.empty_like(x); dw2 = torch.zeros(C, device=dev); db2 = torch.zeros(C, device=dev)
    cs = torch.full((C,), 0.5, device=dev)
    K.layernorm_bwd(dy, x, w, mean, rstd, None, dx2, dw2, db2, dxsum=cs)
    assert torch.allclose(dx2.float(), dxr, atol=3e-2, rtol=3e-2)
    assert torch.allclose(cs, dx2.float().sum(0) + 0.5, atol=2e-3 * math.sqrt(M), rtol=1e-4), (cs - dx2.float().sum(0) - 0.5).abs().max().item()
    assert torch.allclose(dw2, dw, atol=1e-3 * math.sqrt(M), rtol=1e-3) and torch.allclose(db2, db, atol=1e-3 * math.sqrt(M), rtol=1e-3)


def test_bias_grad_and_embedding(dev):
    from trustworthy_dl.ops import nn as K
    torch.manual_seed(0)
    dy = torch.randn(1000, 768, device=dev).bfloat16()
    db = torch.zeros(768, device=dev)
    K.bias_grad(dy, db)
    assert torch.allclose(db, dy.float().sum(0), atol=0.05, rtol=1e-3)
    T, C, V = 128, 256, 1000
    wte = torch.randn(V, C, device=dev).bfloat16(); wpe = torch.randn(T, C, device=dev).bfloat16()
    idx = torch.randint(0, V, (4, T), device=dev)
    out = torch.empty(4 * T, C, device=dev, dtype=torch.bfloat16)
    K.embedding_fwd(idx, wte, wpe, out, T)
    ref = wte[idx].float() + wpe.float()[None]
    assert torch.allclose(out.view(4, T, C).float(), ref, atol=2e-2, rtol=2e-2)
    dx = torch.randn(4 * T, C, device=dev).bfloat16()
    dwte = torch.zeros(V, C, device=dev); dwpe = torch.zeros(T, C, device=dev)
    K.embedding_bwd(idx, dx, dwte, dwpe, T)
    rwte = torch.zeros(V, C, device=dev).index_add_(0, idx.view(-1), dx.float())
    assert torch.allclose(dwte, rwte, atol=1e-3) and torch.allclose(dwpe, dx.float().view(4, T, C).sum(0), atol=1e-3)


@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_adamw(dev, gdtype):
    from trustworthy_dl.ops import nn as K
    torch.manual_seed(0)
    n, nd = 4096, 3000
    p = torch.randn(n, device=dev); g = torch.randn(n, device=dev).to(gdtype)
    m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev); p16 = torch.empty(n, device=dev, dtype=torch.bfloat16)
    pr, mr, vr = p.clone(), m.clone(), v.clone()
    hyper = torch.zeros(4, device=dev)
    for step in (1, 2, 3):
        K.adam_tick(hyper, 1e-2, 0.9, 0.999)
        K.adamw(p, p16, g, m, v, nd, 123.0, 0.9, 0.999, 1e-8, 0.1, 99, hyper=hyper)      # lr/step overridden by hyper
        R.adamw_step(pr[:nd], g[:nd], mr[:nd], vr[:nd], 1e-2, 0.9, 0.999, 1e-8, 0.1, step)
        R.adamw_step(pr[nd:], g[nd:], mr[nd:], vr[nd:], 1e-2, 0.9, 0.999, 1e-8, 0.0, step)
    assert torch.allclose(p, pr, atol=1e-5, rtol=1e-4) and torch.allclose(m, mr, atol=1e-6) and torch.allclose(v, vr, atol=1e-6)
    assert torch.allclose(p16.float(), p, atol=1e-2, rtol=1e-2)


def test_cross_entropy(dev):
    from trustworthy_dl.ops import nn as K
    torch.manual_seed(0)
    M, V, Vp = 64, 50257, 50304
    logits = (torch.randn(M, Vp, device=dev) * 3).bfloat16()
    tgt = torch.randint(0, V, (M,), device=dev)
    ref_loss, ref_d = R.cross_entropy_fwd_bwd(logits, tgt, V)
    row = torch.empty(M, device=dev); tot = torch.zeros(1, device=dev)
    K.xent_fwd_bwd(logits, tgt, row, tot, V, 1.0 / M)
    assert torch.allclose(row, ref_loss, atol=2e-3, rtol=1e-3)
    assert tot.item() == pytest.approx(ref_loss.sum().item(), rel=1e-4)
    assert torch.allclose(logits.float(), ref_d, atol=2e-4 / M * 64, rtol=2e-2)
    assert torch.count_nonzero(logits[:, V:]) == 0


def test_pack_norm_multinorm_checksum_drift_poison(dev):
    from trustworthy_dl.ops import nn as K
    from trustworthy_dl.ops.multi_tensor import multi_tensor_l2norm, per_tensor_l2norms
    from trustworthy_dl.ops.integrity import checksum64, drift_sq_and_update
    from trustworthy_dl.ops.attack_ops import poison_inplace
    torch.manual_seed(0)
    src = torch.randn(1 << 20, device=dev)
    dst = torch.empty(1 << 20, device=dev, dtype=torch.bfloat16); ss = torch.zeros(1, device=dev, dtype=torch.float64)
    K.pack_norm(src, dst, ss)
    assert ss.item() == pytest.approx(src.double().pow(2).sum().item(), rel=1e-6)
    assert torch.equal(dst, src.bfloat16())
    ss.zero_(); K.pack_norm(src, dst, ss, torch.full((1,), -3.0, device=dev))
    assert torch.equal(dst, (src * -3.0).bfloat16()) and ss.item() == pytest.approx(9 * src.double().pow(2).sum().item(), rel=1e-6)
    ts = [torch.randn(s, device=dev) for s in (7, 1000, 12345, 1 << 16)]
    norms = per_tensor_l2norms(ts)
    assert torch.allclose(norms, torch.stack([t.norm() for t in ts]), rtol=1e-4)
    assert multi_tensor_l2norm(ts).item() == pytest.approx(torch.cat(ts).norm().item(), rel=1e-4)
    x = torch.randn(100_003, device=dev)
    assert checksum64(x) == R.fletcher64(x.cpu())
    y = x.clone(); y[5000] += 1e-3
    assert checksum64(y) != checksum64(x)
    xb = torch.randn(4096, device=dev).bfloat16()
    assert checksum64(xb) == R.fletcher64(xb.cpu())
    prev = x.clone(); cur = x + 0.01
    d = drift_sq_and_update(cur, prev)
    assert d == pytest.approx(1e-4 * x.numel(), rel=1e-2) and torch.equal(prev, cur)
    t = torch.randn(1 << 16, device=dev); t0 = t.clone()
    poison_inplace(t, "scale", 0.5, 1); assert torch.allclose(t, t0 * 11.0)
    t = t0.clone(); poison_inplace(t, "sign_flip", 0.5, 1); assert torch.allclose(t, t0 * -3.0)
    t = t0.clone(); poison_inplace(t, "noise", 0.5, 7)
    noise = (t - t0)
    assert noise.std().item() == pytest.approx(2.0 * t0.pow(2).mean().sqrt().item(), rel=0.05) and abs(noise.mean().item()) < 0.05


@pytest.mark.parametrize("W,dtype", [(1, torch.bfloat16), (2, torch.bfloat16), (3, torch.float32), (4, torch.bfloat16),
                                     (8, torch.bfloat16), (8, torch.float32)])
def test_trust_reduce_emulated_ranks(dev, W, dtype):
    """All ranks' buckets on one GPU; the math kernel is launched once per emulated rank (no waits)."""
    from trustworthy_dl.ops import _lib
    from trustworthy_dl.ops._lib import i32, i64, ptr, stream_ptr
    torch.manual_seed(W)
    vec = 8 if dtype == torch.bfloat16 else 4
    N = vec * W * 1000
    base = torch.randn(N)
    bufs = [(base + 0.5 * torch.randn(N)).to(dtype).to(dev) for _ in range(W)]
    if W >= 3:
        bufs[1] = (-3.0 * bufs[1].float()).to(dtype)              # a sign-flipped, scaled rank
    orig = [b.clone() for b in bufs]
    coeff = torch.rand(W, device=dev); coeff = coeff / coeff.sum()
    outs = [torch.empty_like(b) for b in bufs]
    partial = torch.zeros(W + 1, dtype=torch.float64, device=dev)
    in_arr = (ctypes.c_void_p * W)(*[b.data_ptr() for b in bufs])
    out_arr = (ctypes.c_void_p * W)(*[o.data_ptr() for o in outs])
    per = N // vec // W
    for r in range(W):
        _lib.call("tdl_trust_reduce", in_arr, out_arr, i32(W), i32(_lib.DTYPE_CODES[dtype]), i64(r * per), i64((r + 1) * per),
                  ptr(coeff), ptr(partial), stream_ptr())
    red, stats = R.trust_weighted_reduce([o.float() for o in orig], coeff.cpu().to(dev))
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    for o in outs:
        assert torch.allclose(o.float(), red, atol=tol, rtol=tol)
    p = partial.cpu()
    assert torch.allclose(p[:W].float(), stats[:, 1].cpu(), rtol=1e-3, atol=1e-2)
    assert p[W].item() == pytest.approx(stats[0, 2].item(), rel=1e-3)
    if W >= 3:
        cos = p[:W] / (torch.stack([o.float().norm() for o in orig]).cpu().double() * p[W].sqrt())
        assert cos[1] < 0 < cos[0]


def test_device_trust_state_machine_matches_host(dev):
    """Single-rank exchange -> coefficients -> reduce -> finalize; device trust follows the host TrustManager."""
    from trustworthy_dl import TrustManager
    from trustworthy_dl.parallel.symm import SymmetricTrustReducer
    r = SymmetricTrustReducer(8 * 1024, torch.bfloat16, dev)
    tm = TrustManager(1)
    for step in range(5):
        r.buffer.copy_(torch.randn(8 * 1024, device=dev).bfloat16())
        before = r.buffer.clone()
        r.enqueue_local_norm()
        r.enqueue_reduce(update_trust=True)
        torch.cuda.synchronize()
        tm.update_trust_score(0, 0.0, 1.0)
        assert r.trust.item() == pytest.approx(tm.get_trust_score(0), abs=1e-5)
        assert r.coeff.item() == pytest.approx(1.0) and torch.equal(r.buffer, before)
        assert r.stats[0, 0].item() == pytest.approx(before.float().norm().item(), rel=1e-3)
        assert r.stats[0, 3].item() == pytest.approx(1.0, abs=1e-3) and r.flags.item() == 0


def test_param_gather_checksum_and_drift(dev):
    from trustworthy_dl.parallel.param_gather import gather_shards
    W, n = 4, 8 * 4001
    torch.manual_seed(0)
    shards = [torch.randn(n, device=dev).bfloat16() for _ in range(W)]
    dst = torch.zeros(W * n, device=dev, dtype=torch.bfloat16)
    prev = torch.cat(shards).clone()
    prev[n:2 * n] += 0.25                                     # shard 1 drifted
    checks, drift = gather_shards([s.data_ptr() for s in shards], W, torch.bfloat16, dst, prev, n)
    assert torch.equal(dst, torch.cat(shards)) and torch.equal(prev, dst)
    for i in range(W):
        assert checks[i] == R.fletcher64(shards[i].cpu())
    d = drift.sqrt().cpu()
    assert d[0] == 0 and d[2] == 0 and d[1] > 10
    shards[2][100] += 1.0                                     # tamper after "publication"
    checks2, _ = gather_shards([s.data_ptr() for s in shards], W, torch.bfloat16, dst, None, n)
    assert checks2[2] != checks[2] and checks2[0] == checks[0]


@pytest.mark.parametrize("channels_last", [False, True])
def test_conv_epilogue(dev, channels_last):
    from trustworthy_dl.ops.conv_epilogue import fused_affine_add_relu, fused_bias_relu
    torch.manual_seed(0)
    x = torch.randn(4, 16, 8, 8, device=dev, requires_grad=True)
    res = torch.randn(4, 16, 8, 8, device=dev, requires_grad=True)
    if channels_last:
        x = x.detach().contiguous(memory_format=torch.channels_last).requires_grad_(True)
        res = res.detach().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    scale = (torch.rand(16, device=dev) + 0.5).requires_grad_(True)
    shift = torch.randn(16, device=dev, requires_grad=True)
    y = fused_affine_add_relu(x, scale, shift, res)
    ref = torch.relu(x * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1) + res)
    assert torch.allclose(y, ref, atol=1e-5)
    g = torch.randn_like(ref)
    got = torch.autograd.grad(y, (x, scale, shift, res), g)
    exp = torch.autograd.grad(ref, (x, scale, shift, res), g)
    for a, b in zip(got, exp):
        assert torch.allclose(a, b, atol=1e-3, rtol=1e-3)
    xb = torch.randn(2, 8, 4, 4, device=dev).bfloat16()
    b = torch.randn(8, device=dev)
    assert torch.allclose(fused_bias_relu(xb, b).float(), torch.relu(xb.float() + b.view(1, -1, 1, 1)), atol=2e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_resnet_bnact_matches_batchnorm_add_relu(dev, dtype):
    """Training-mode BatchNorm + residual + ReLU through the K13 epilogue (models/resnet.py:BNAct) against the stock
    three-op sequence: outputs, input / residual / affine gradients and running statistics."""
    from trustworthy_dl.models.resnet import BNAct
    torch.manual_seed(0)
    ours, ref = BNAct(32).to(dev), torch.nn.BatchNorm2d(32).to(dev)
    with torch.no_grad():
        ours.weight.uniform_(0.5, 1.5); ours.bias.normal_()
    ref.load_state_dict(ours.state_dict())
    x = (torch.randn(8, 32, 14, 14, device=dev) * 2 + 0.5).to(dtype)
    r = torch.randn(8, 32, 14, 14, device=dev).to(dtype)
    x1, r1 = x.clone().requires_grad_(True), r.clone().requires_grad_(True)
    x2, r2 = x.float().clone().requires_grad_(True), r.float().clone().requires_grad_(True)
    y = ours(x1, residual=r1, relu=True)
    y_ref = torch.relu(ref(x2) + r2)
    tol = 1e-4 if dtype == torch.float32 else 3e-2
    assert torch.allclose(y.float(), y_ref, atol=tol, rtol=tol)
    g = torch.randn_like(y_ref)
    y.backward(g.to(dtype)); y_ref.backward(g)
    gt = 2e-3 if dtype == torch.float32 else 6e-2
    assert torch.allclose(x1.grad.float(), x2.grad, atol=gt, rtol=gt), (x1.grad.float() - x2.grad).abs().max().item()
    assert torch.allclose(r1.grad.float(), r2.grad, atol=gt, rtol=gt)
    assert torch.allclose(ours.weight.grad, ref.weight.grad, atol=gt * 30, rtol=gt)
    assert torch.allclose(ours.bias.grad, ref.bias.grad, atol=gt * 30, rtol=gt)
    assert torch.allclose(ours.running_mean, ref.running_mean, atol=1e-2) and torch.allclose(ours.running_var, ref.running_var, rtol=2e-2)
    ours.eval(); ref.eval()
    assert torch.allclose(ours(x, relu=False).float(), ref(x.float()), atol=tol * 2, rtol=tol * 2)
