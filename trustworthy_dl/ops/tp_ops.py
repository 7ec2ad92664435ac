"""Tensor-parallel GEMMs fused with their collectives over peer-mapped memory (K7, K8).

* :func:`allgather_gemm` — column-parallel linear on a sequence-sharded input: every rank holds
  ``[M/W, K]`` rows; the GEMM's TMA producer reads each A tile from the owning peer as it is needed, so the
  "all-gather" is spread over the whole main loop instead of preceding it.
* :func:`gemm_scatter` + :func:`reduce_slots` — row-parallel linear: each rank's partial ``[M, N]`` product
  is stored tile by tile from the epilogue into the owner's staging slots; after the barrier the owner sums
  its W slots (fused with the residual add).
Callers provide the peer pointer lists (from :class:`trustworthy_dl.parallel.symm.SymmetricBuffer`) and place a
cross-rank barrier (``tdl_exchange``) between producer and consumer kernels.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import _lib
from ._lib import i32, i64, ptr, stream_ptr


def _arr(ptrs: List[int]):
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


def allgather_gemm(a_shard_ptrs: List[int], rank: int, shard_rows: int, w: torch.Tensor, out: torch.Tensor, *,
                   bias: Optional[torch.Tensor] = None, epilogue: int = 0, aux_in=None, aux_out=None, block_n: int = 0,
                   b_mn: bool = False, colsum: Optional[torch.Tensor] = None):
    """out [W*shard_rows, N] = AG(a shards) @ w^T (``w`` [N, K]) or @ w (``b_mn``: ``w`` [K, N], the dgrad of a row-parallel
    layer); ``colsum`` (fp32 [N]) += column sums of the bf16 output."""
    W = len(a_shard_ptrs)
    if b_mn:
        K, N = w.shape
    else:
        N, K = w.shape
    bias_f32 = 1 if (bias is not None and bias.dtype == torch.float32) else 0
    _lib.call("tdl_allgather_gemm_bf16_ex", _arr(a_shard_ptrs), i32(W), i32(rank), i64(shard_rows), ptr(w), ptr(out), i64(N), i64(K),
              i64(K), i64(w.shape[1]), i64(N), i32(b_mn), ptr(bias), i32(bias_f32), i32(epilogue), ptr(aux_in), ptr(aux_out), i64(N),
              i32(block_n), ptr(colsum), stream_ptr())


def gemm_scatter(a: torch.Tensor, w: torch.Tensor, staging_ptrs: List[int], rank: int, shard_rows: int, *,
                 b_mn: bool = False, block_n: int = 0):
    """a [M, K] @ w[N, K]^T (or w [K, N] when b_mn) -> partial rows scattered to the owners' slot `rank`."""
    W = len(staging_ptrs)
    M, K = a.shape
    N = w.shape[1] if b_mn else w.shape[0]
    assert M == shard_rows * W
    _lib.call("tdl_gemm_scatter_bf16", ptr(a), ptr(w), _arr(staging_ptrs), i32(W), i32(rank), i64(shard_rows), i64(N), i64(K),
              i64(K), i64(w.shape[1]), i64(N), i32(b_mn), i32(block_n), stream_ptr())


def reduce_slots(staging: torch.Tensor, W: int, out: torch.Tensor, residual: Optional[torch.Tensor] = None,
                 bias: Optional[torch.Tensor] = None):
    """out [rows, cols] = sum of the W staged partial slots (+ fp32 bias[cols]) (+ residual)."""
    if bias is None:
        _lib.call("tdl_reduce_slots", ptr(staging), i32(W), i64(out.numel()), ptr(residual), ptr(out), stream_ptr())
    else:
        rows, cols = out.shape
        _lib.call("tdl_reduce_slots_bias", ptr(staging), i32(W), i64(rows), i64(cols), ptr(bias), ptr(residual), ptr(out),
                  stream_ptr())


_gather_dummy = {}


def gather_rows(shard_ptrs: List[int], shard_elems: int, dst: torch.Tensor):
    """dst [W * shard_elems] <- the W peer shards (plain P2P pull with the K9 kernel, integrity outputs discarded)."""
    W = len(shard_ptrs)
    key = dst.device.index
    if key not in _gather_dummy:
        _gather_dummy[key] = (torch.zeros(16, dtype=torch.int64, device=dst.device), torch.zeros(8, dtype=torch.float64, device=dst.device))
    sums, drift = _gather_dummy[key]
    _lib.call("tdl_param_gather", _arr(shard_ptrs), i32(W), i32(_lib.DTYPE_CODES[dst.dtype]), ptr(dst), None, i64(shard_elems),
              ptr(sums), ptr(drift), stream_ptr())


def local_gemm(a: torch.Tensor, w: torch.Tensor, out: torch.Tensor, *, bias: Optional[torch.Tensor] = None, epilogue: int = 0,
               aux_in=None, aux_out=None, block_n: int = 0, b_mn: bool = False, colsum: Optional[torch.Tensor] = None):
    """Plain tcgen05 GEMM on an already gathered operand, same keyword surface as the fused forms."""
    from . import nn as K
    M = a.shape[0]
    if b_mn:
        Kd, N = w.shape
    else:
        N, Kd = w.shape
    K.gemm(a, w, out, M, N, Kd, lda=Kd, ldb=w.shape[1], ldc=N, b_mn=b_mn, bias=bias, epilogue=epilogue, aux_in=aux_in,
           aux_out=aux_out, ld_aux=N, block_n=block_n, colsum=colsum)
