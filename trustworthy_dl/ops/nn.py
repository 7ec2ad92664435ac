"""Thin ctypes wrappers over the sm_100a kernels used by the fused GPT-2 engine (raw pointers,
current stream, no allocation — everything here is CUDA-graph capturable)."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib
from ._lib import f32, i32, i64, ptr, stream_ptr

EPI_NONE, EPI_GELU, EPI_RESIDUAL, EPI_DGELU = 0, 1, 2, 3
EPI_GELU_DG, EPI_MUL = 4, 5      # forward stores gelu'(pre) in aux_out; backward is then out = acc * aux_in


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, *, lda: int, ldb: int,
         ldc: int, a_mn: bool = False, b_mn: bool = False, bias: Optional[torch.Tensor] = None, epilogue: int = EPI_NONE,
         aux_in: Optional[torch.Tensor] = None, aux_out: Optional[torch.Tensor] = None, ld_aux: int = 0,
         atomic_f32: bool = False, split_k: int = 1, block_n: int = 0, colsum: Optional[torch.Tensor] = None) -> None:
    """out[M,N] (+)= A @ B^T on tcgen05 (see include/tdl_kernels.h for the layout flags).  ``colsum`` (fp32 [N]):
    += column sums of the bf16 output, fused into the epilogue."""
    bias_f32 = 1 if (bias is not None and bias.dtype == torch.float32) else 0
    if colsum is not None:
        _lib.call("tdl_gemm_bf16_colsum", ptr(a), ptr(b), ptr(out), i64(M), i64(N), i64(K), i64(lda), i64(ldb), i64(ldc),
                  i32(a_mn), i32(b_mn), ptr(bias), i32(bias_f32), i32(epilogue), ptr(aux_in), ptr(aux_out),
                  i64(ld_aux or ldc), i32(block_n), ptr(colsum), stream_ptr())
        return
    _lib.call("tdl_gemm_bf16", ptr(a), ptr(b), ptr(out), i64(M), i64(N), i64(K), i64(lda), i64(ldb), i64(ldc),
              i32(a_mn), i32(b_mn), ptr(bias), i32(bias_f32), i32(epilogue), ptr(aux_in), ptr(aux_out),
              i64(ld_aux or ldc), i32(atomic_f32), i32(split_k), i32(block_n), stream_ptr())


def linear_fwd(x, w, out, bias=None, epilogue=EPI_NONE, aux_in=None, aux_out=None):
    """x [M,K] @ w[N,K]^T (+bias) -> out [M,N]; all row-major contiguous."""
    M, K = x.shape
    N = w.shape[0]
    gemm(x, w, out, M, N, K, lda=K, ldb=K, ldc=N, bias=bias, epilogue=epilogue, aux_in=aux_in, aux_out=aux_out, ld_aux=N)


def linear_dgrad(dy, w, dx, epilogue=EPI_NONE, aux_in=None, colsum=None):
    """dx [M,K] = dy [M,N] @ w [N,K]  (w consumed MN-major, no transpose copy); colsum [K] fp32 += column sums of dx."""
    M, N = dy.shape
    K = w.shape[1]
    gemm(dy, w, dx, M, K, N, lda=N, ldb=K, ldc=K, b_mn=True, epilogue=epilogue, aux_in=aux_in, ld_aux=K, colsum=colsum)


def linear_wgrad(dy, x, dw32, split_k=1):
    """dw32 [N,K] (fp32) += dy [M,N]^T @ x [M,K]  (both consumed MN-major)."""
    M, N = dy.shape
    K = x.shape[1]
    gemm(dy, x, dw32, N, K, M, lda=N, ldb=K, ldc=K, a_mn=True, b_mn=True, atomic_f32=True, split_k=split_k)


def attention_fwd(qkv, out, lse, B, T, H):
    _lib.call("tdl_attention_fwd", ptr(qkv), ptr(out), ptr(lse), i64(B), i64(T), i64(H), i64(64), stream_ptr())


def attention_bwd(qkv, out, d_out, lse, delta, dqkv, B, T, H, dbias=None):
    """dbias (fp32 [3C], optional) += column sums of dqkv (the qkv bias gradient), fused into the accumulator write-out."""
    _lib.call("tdl_attention_bwd_dbias", ptr(qkv), ptr(out), ptr(d_out), ptr(lse), ptr(delta), ptr(dqkv), ptr(dbias),
              i64(B), i64(T), i64(H), i64(64), stream_ptr())


def layernorm_fwd(x, w, b, y, mean, rstd, eps=1e-5):
    M, C = x.shape
    _lib.call("tdl_layernorm_fwd", ptr(x), ptr(w), ptr(b), ptr(y), ptr(mean), ptr(rstd), i64(M), i64(C), f32(eps), stream_ptr())


def layernorm_bwd(dy, x, w, mean, rstd, dres, dx, dw, db, dxsum=None):
    """dx = dres + LN'(dy); dw/db += dgamma/dbeta; dxsum (fp32 [C], optional) += column sums of dx."""
    M, C = x.shape
    _lib.call("tdl_layernorm_bwd_colsum", ptr(dy), ptr(x), ptr(w), ptr(mean), ptr(rstd), ptr(dres), ptr(dx), ptr(dw), ptr(db),
              ptr(dxsum), i64(M), i64(C), stream_ptr())


def bias_grad(dy, db32):
    M, N = dy.shape
    _lib.call("tdl_bias_grad", ptr(dy), ptr(db32), i64(M), i64(N), i64(N), stream_ptr())


def embedding_fwd(idx, wte, wpe, out, T):
    BT, C = out.shape
    _lib.call("tdl_embedding_fwd", ptr(idx), ptr(wte), ptr(wpe), ptr(out), i64(BT), i64(T), i64(C), stream_ptr())


def embedding_bwd(idx, dx, dwte32, dwpe32, T):
    BT, C = dx.shape
    _lib.call("tdl_embedding_bwd", ptr(idx), ptr(dx), ptr(dwte32), ptr(dwpe32), i64(BT), i64(T), i64(C), stream_ptr())


def adamw(p32, p16, grad, m, v, n_decay, lr, beta1, beta2, eps, wd, step, grad_scale=None, hyper=None):
    _lib.call("tdl_adamw", ptr(p32), ptr(p16), ptr(grad), i32(_lib.dtype_code(grad)), ptr(m), ptr(v), i64(p32.numel()),
              i64(n_decay), f32(lr), f32(beta1), f32(beta2), f32(eps), f32(wd), i64(step), ptr(grad_scale), ptr(hyper),
              stream_ptr())


def adam_tick(hyper, lr, beta1, beta2):
    _lib.call("tdl_adam_tick", ptr(hyper), f32(lr), f32(beta1), f32(beta2), stream_ptr())


def xent_fwd_bwd(logits, targets, row_loss, loss_sum, V, grad_scale):
    M, Vp = logits.shape
    _lib.call("tdl_xent_fwd_bwd", ptr(logits), ptr(targets), ptr(row_loss), ptr(loss_sum), i64(M), i64(V), i64(Vp),
              f32(grad_scale), stream_ptr())


def pack_norm(src32, dst16, sumsq, scale=None):
    _lib.call("tdl_pack_norm", ptr(src32), ptr(dst16), i64(src32.numel()), ptr(sumsq), ptr(scale), stream_ptr())


def cast_f32_bf16(src, dst):
    _lib.call("tdl_cast_f32_bf16", ptr(src), ptr(dst), i64(src.numel()), stream_ptr())


def cast_bf16_f32(src, dst):
    _lib.call("tdl_cast_bf16_f32", ptr(src), ptr(dst), i64(src.numel()), stream_ptr())
