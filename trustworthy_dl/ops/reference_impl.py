"""Plain-PyTorch restatements of every fused op.

Two jobs: (1) they are the CPU / gloo execution path, so the whole framework runs without a
GPU; (2) they are the fp32 oracles the sm_100a kernels in ``csrc/`` are tested against
(tests/test_gpu_*.py).  Nothing here is on the hot path when a B200 is present.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

STAT_NAMES = ("mean", "std", "min", "max", "median", "skewness", "kurtosis",
              "percentile_25", "percentile_75", "norm_l1", "norm_l2", "norm_inf")
NUM_STATS = len(STAT_NAMES)


def _interp_quantile(sorted_x: torch.Tensor, q: float) -> torch.Tensor:
    """numpy 'linear' percentile on an already sorted 1-D tensor."""
    n = sorted_x.numel()
    pos = q * (n - 1)
    lo = int(math.floor(pos))
    hi = min(lo + 1, n - 1)
    frac = pos - lo
    return sorted_x[lo] + (sorted_x[hi] - sorted_x[lo]) * frac


def tensor_statistics(x: torch.Tensor) -> torch.Tensor:
    """The detector's 12 statistics (reference attack_detector.py:185-200) as a float64
    vector in ``STAT_NAMES`` order: population std (ddof=0), biased skewness, biased Fisher
    kurtosis, linearly interpolated percentiles."""
    v = x.detach().reshape(-1).to(torch.float64)
    n = v.numel()
    mean = v.mean()
    c = v - mean
    m2 = (c * c).mean()
    m3 = (c * c * c).mean()
    m4 = (c * c * c * c).mean()
    std = m2.sqrt()
    # scipy returns nan for constant input; keep that behaviour (nan never passes std>0 filter)
    skew = m3 / m2.pow(1.5)
    kurt = m4 / (m2 * m2) - 3.0
    s, _ = torch.sort(v)
    out = torch.stack([
        mean, std, s[0], s[-1], _interp_quantile(s, 0.5), skew, kurt,
        _interp_quantile(s, 0.25), _interp_quantile(s, 0.75),
        v.abs().sum(), (v * v).sum().sqrt(), v.abs().max(),
    ])
    assert out.numel() == NUM_STATS and n > 0
    return out


def moments_raw(x: torch.Tensor) -> torch.Tensor:
    """[sum, sum2, sum3, sum4, min, max, sum|x|, max|x|] in float64 (oracle for the
    single-pass moment kernel)."""
    v = x.detach().reshape(-1).to(torch.float64)
    return torch.stack([v.sum(), (v * v).sum(), (v ** 3).sum(), (v ** 4).sum(), v.min(), v.max(),
                        v.abs().sum(), v.abs().max()])


# ----------------------------------------------------------------------------- NN ops
def layernorm_fwd(x, w, b, eps=1e-5):
    xf = x.float()
    mean = xf.mean(-1, keepdim=True)
    var = xf.var(-1, unbiased=False, keepdim=True)
    rstd = (var + eps).rsqrt()
    y = (xf - mean) * rstd * w.float() + b.float()
    return y.to(x.dtype), mean.squeeze(-1), rstd.squeeze(-1)


def layernorm_bwd(dy, x, w, mean, rstd):
    """Returns (dx, dw, db) in fp32."""
    xf, dyf, wf = x.float(), dy.float(), w.float()
    xhat = (xf - mean.unsqueeze(-1)) * rstd.unsqueeze(-1)
    dnorm = dyf * wf
    n = x.shape[-1]
    dx = (dnorm - dnorm.mean(-1, keepdim=True) - xhat * (dnorm * xhat).mean(-1, keepdim=True)) * rstd.unsqueeze(-1)
    dw = (dyf * xhat).reshape(-1, n).sum(0)
    db = dyf.reshape(-1, n).sum(0)
    return dx, dw, db


_GELU_C = math.sqrt(2.0 / math.pi)


def gelu_tanh(x):
    xf = x.float()
    return 0.5 * xf * (1.0 + torch.tanh(_GELU_C * (xf + 0.044715 * xf ** 3)))


def gelu_tanh_bwd(x, dy):
    xf = x.float()
    u = _GELU_C * (xf + 0.044715 * xf ** 3)
    t = torch.tanh(u)
    du = _GELU_C * (1.0 + 3 * 0.044715 * xf * xf)
    return dy.float() * (0.5 * (1.0 + t) + 0.5 * xf * (1.0 - t * t) * du)


def causal_attention_fwd(q, k, v, scale=None):
    """q,k,v: [B, H, T, D] -> (o [B,H,T,D] fp32, lse [B,H,T] fp32)."""
    qf, kf, vf = q.float(), k.float(), v.float()
    T = q.shape[-2]
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    s = (qf @ kf.transpose(-1, -2)) * scale
    mask = torch.ones(T, T, dtype=torch.bool, device=q.device).tril()
    s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse.unsqueeze(-1))
    return p @ vf, lse


def causal_attention_bwd(q, k, v, o, lse, do, scale=None):
    qf, kf, vf, of, dof = q.float(), k.float(), v.float(), o.float(), do.float()
    T = q.shape[-2]
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    s = (qf @ kf.transpose(-1, -2)) * scale
    mask = torch.ones(T, T, dtype=torch.bool, device=q.device).tril()
    p = torch.exp(s - lse.unsqueeze(-1)).masked_fill(~mask, 0.0)
    dv = p.transpose(-1, -2) @ dof
    dp = dof @ vf.transpose(-1, -2)
    delta = (dof * of).sum(-1, keepdim=True)
    ds = p * (dp - delta) * scale
    dq = ds @ kf
    dk = ds.transpose(-1, -2) @ qf
    return dq, dk, dv


def cross_entropy_fwd_bwd(logits, targets, vocab: int, grad_scale: Optional[float] = None):
    """Mean CE over rows using only the first ``vocab`` columns (padding columns get zero
    gradient). Returns (per-row loss fp32 [M], dlogits fp32 [M, Vp])."""
    M, Vp = logits.shape
    lf = logits.float()[:, :vocab]
    lse = torch.logsumexp(lf, dim=-1)
    loss = lse - lf.gather(1, targets.view(-1, 1).long()).squeeze(1)
    p = torch.exp(lf - lse.unsqueeze(1))
    p.scatter_add_(1, targets.view(-1, 1).long(), -torch.ones(M, 1, device=logits.device))
    gs = (1.0 / M) if grad_scale is None else grad_scale
    d = torch.zeros(M, Vp, device=logits.device)
    d[:, :vocab] = p * gs
    return loss, d


def adamw_step(p32, g, m, v, lr, beta1, beta2, eps, wd, step, grad_scale=1.0):
    """In-place AdamW on fp32 master params; returns nothing. ``g`` may be bf16."""
    gf = g.float() * grad_scale
    m.mul_(beta1).add_(gf, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(gf, gf, value=1 - beta2)
    mhat = m / (1 - beta1 ** step)
    vhat = v / (1 - beta2 ** step)
    p32.sub_(lr * (mhat / (vhat.sqrt() + eps) + wd * p32))


# -------------------------------------------------------- trust-weighted reduction
def median_of_ranks(stack: torch.Tensor) -> torch.Tensor:
    """Coordinate-wise median over dim 0 of ``[W, N]``.  For even W this is the mean of the two
    middle values (numpy convention; what the 8-input sorting network in the kernel produces)."""
    W = stack.shape[0]
    s, _ = torch.sort(stack.float(), dim=0)
    if W % 2:
        return s[W // 2]
    return 0.5 * (s[W // 2 - 1] + s[W // 2])


def trust_weighted_reduce(bucket_per_rank: Sequence[torch.Tensor], weights: torch.Tensor,
                          clip_norm: float = 0.0, norms: Optional[torch.Tensor] = None
                          ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Oracle of the fused trust-weighted all-reduce (kernel K2).

    ``bucket_per_rank``: W tensors of N gradient elements.  ``weights``: [W] normalised trust
    weights.  ``clip_norm`` > 0 clips rank i by ``min(1, clip_norm / ||g_i||)`` before weighting.
    Returns ``(reduced [N] fp32, stats [W, 4] fp32)`` with stats columns
    ``(l2_norm, dot(g_i, median), ||median||^2, cosine(g_i, median))``.
    """
    g = torch.stack([b.reshape(-1).float() for b in bucket_per_rank])       # [W, N]
    W = g.shape[0]
    if norms is None:
        norms = g.norm(dim=1)
    # non-finite elements are read as 0 (as in the kernel): a rank shipping NaN/Inf has coefficient 0 through its
    # non-finite norm, but 0 * NaN would still poison the aggregate and the median
    g = torch.nan_to_num(g, nan=0.0, posinf=0.0, neginf=0.0)
    med = median_of_ranks(g)
    dots = g @ med
    med2 = (med * med).sum()
    cos = dots / (norms * med2.sqrt()).clamp_min(1e-30)
    clip = torch.ones(W, device=g.device)
    if clip_norm > 0:
        clip = torch.clamp(clip_norm / norms.clamp_min(1e-30), max=1.0)
    coeff = weights.float().to(g.device) * clip
    reduced = (coeff.unsqueeze(1) * g).sum(0)
    stats = torch.stack([norms, dots, med2.expand(W), cos], dim=1)
    return reduced, stats


def fletcher64(x: torch.Tensor) -> int:
    """Order-sensitive 64-bit checksum over the raw 32-bit words of ``x`` (oracle of the
    parameter-integrity kernel): two running sums modulo 2^32, packed (hi=sum2, lo=sum1)."""
    raw = x.detach().contiguous().view(torch.uint8).cpu().numpy()
    import numpy as np
    pad = (-raw.size) % 4
    if pad:
        raw = np.concatenate([raw, np.zeros(pad, dtype=np.uint8)])
    words = raw.view(np.uint32).astype(np.uint64)
    n = words.size
    s1 = int(words.sum() % (1 << 32))
    # sum2 = sum_{i} (n - i) * w_i  mod 2^32  (closed form of the running sum of s1)
    coef = (np.arange(n, 0, -1, dtype=np.uint64)) % (1 << 32)
    s2 = int(((coef * words) % (1 << 32)).sum() % (1 << 32))
    return (s2 << 32) | s1


def robust_coefficients(norms: torch.Tensor, trust_weights: torch.Tensor, clip_ratio: float = 2.0,
                        norm_ratio_max: float = 2.5) -> torch.Tensor:
    """Per-rank reduction coefficients (oracle of the device-side weight kernel, K3):
    ``c_i = w_i * valid_i * min(1, clip_ratio*med/norm_i) / sum_j(w_j * valid_j)`` where ``med`` is the
    (lower) median bucket norm over ranks with non-zero trust weight and ``valid_i`` rejects non-finite or
    ``> norm_ratio_max * med`` norms *in the same step* (the cosine check lags one step)."""
    n = norms.double()
    w = trust_weights.double().to(n.device)
    finite = torch.isfinite(n)
    live = (w > 0) & finite
    if not bool(live.any()):
        return torch.zeros_like(n, dtype=torch.float32)
    vals, _ = torch.sort(n[live])
    k = vals.numel()
    med = vals[(k - 1) // 2]          # lower median: one exploding norm cannot drag the reference up (k = 2!)
    valid = live & ((n <= norm_ratio_max * med) | (med <= 0))
    clip = torch.ones_like(n)
    if clip_ratio > 0 and float(med) > 0:
        clip = torch.clamp(clip_ratio * med / n.clamp_min(1e-30), max=1.0)
    wv = torch.where(valid, w, torch.zeros_like(w))
    tot = wv.sum()
    if float(tot) <= 0:
        return torch.zeros_like(n, dtype=torch.float32)
    return torch.where(valid, wv * clip / tot, torch.zeros_like(n)).float()


def bucketed_trust_reduce(grads_per_rank: Sequence[torch.Tensor], bounds: Sequence[Tuple[int, int]], trust_weights: torch.Tensor,
                          clip_ratio: float = 2.0, norm_ratio_max: float = 2.5, cosine_min: float = 0.0):
    """Oracle of one STEP of the bucketed aggregation (engine/gpt2_engine.py): every bucket ``[a, b)`` is reduced with
    coefficients from that bucket's own norms (K1 -> K3 coefficients -> K2); the verdict is taken once per step from the
    accumulated squared norms, the "rejected in some bucket" marks and the accumulated dot / |median|^2 sums, exactly as
    ``finalize_trust_kernel`` does.

    Returns ``(reduced [N], info)`` with ``info = {norms, cos, flags, rejected}`` (per rank)."""
    W = len(grads_per_rank)
    g = [x.reshape(-1).float() for x in grads_per_rank]
    out = torch.zeros_like(g[0])
    acc_sq = torch.zeros(W, dtype=torch.float64)
    dots = torch.zeros(W, dtype=torch.float64)
    med2 = 0.0
    rejected = torch.zeros(W, dtype=torch.bool)
    for a, b in bounds:
        parts = [x[a:b] for x in g]
        norms = torch.stack([p.norm() for p in parts])
        acc_sq += norms.double() ** 2
        coeff = robust_coefficients(norms, trust_weights, clip_ratio, norm_ratio_max)
        live = (trust_weights > 0) & torch.isfinite(norms)
        rejected |= live & (coeff == 0)
        red, stats = trust_weighted_reduce(parts, coeff, 0.0, norms)
        out[a:b] = red
        dots += stats[:, 1].double()
        med2 += float(stats[0, 2])
    norms = acc_sq.sqrt()
    mn = med2 ** 0.5
    cos = torch.where((norms > 0) & (mn > 0), dots / (norms * mn).clamp_min(1e-300), torch.ones_like(dots))
    fin = torch.sort(norms[torch.isfinite(norms)]).values
    ref = float(fin[(fin.numel() - 1) // 2]) if fin.numel() else 0.0
    ok = torch.isfinite(norms) & torch.isfinite(cos) & ~rejected
    if W >= 3:
        ok &= cos >= cosine_min
    if W >= 2 and ref > 0:
        ok &= norms <= norm_ratio_max * ref
    return out, {"norms": norms, "cos": cos, "flags": (~ok).int(), "rejected": rejected}
