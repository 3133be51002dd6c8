"""`torch.library` registration of the engine's tensor-level entry points.

The reference wraps `RqVae.forward` in `@torch.compile(mode="reduce-overhead")` (modules/rqvae.py:140).  Dynamo cannot
trace a `ctypes` call, so the calls `functional.py` makes into `librqengine.so` are registered here as custom ops
(`torch.ops.rq_engine.*`) with fake (meta) implementations for shape propagation and an autograd formula: inside a
compiled region the fused quantiser is then ONE opaque node — one launch forward, one launch backward — that
Inductor schedules around and CUDA-graph capture records like any other kernel.  The eager path
(`functional.ResidualQuantizeFn`) is unchanged and stays the default outside `torch.compile`.

The ops only exist for CUDA tensors: there is no CPU kernel behind them (a CPU tensor raises in `_lib.require_cuda`).
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
from torch import Tensor

from . import functional as RF

_NS = "rq_engine"


# ---- fused forward ------------------------------------------------------------------------------
@torch.library.custom_op(f"{_NS}::quantize_forward", mutates_args=(), device_types="cuda")
def quantize_forward(x: Tensor, codebooks: Tensor, mode: int, beta: float, temperature: float, distance: int,
                     gumbel_uniform: Optional[Tensor], seed: int, offset: int,
                     impl: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    emb, res, ids, loss = RF.quantize_forward(x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed,
                                              offset, True, True, True, impl)
    return emb, res, ids, loss


@quantize_forward.register_fake
def _(x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed, offset, impl):
    B, D = x.shape
    Lv = codebooks.shape[0]
    f32 = dict(device=x.device, dtype=torch.float32)
    return (torch.empty((Lv, B, D), **f32), torch.empty((Lv, B, D), **f32),
            torch.empty((Lv, B), device=x.device, dtype=torch.int64), torch.empty((B,), **f32))


# ---- fused backward -----------------------------------------------------------------------------
@torch.library.custom_op(f"{_NS}::quantize_backward", mutates_args=(), device_types="cuda")
def quantize_backward(g_emb: Optional[Tensor], g_res: Optional[Tensor], g_loss: Optional[Tensor], residuals: Tensor,
                      ids: Tensor, codebooks: Tensor, mode: int, beta: float, temperature: float,
                      gumbel_uniform: Optional[Tensor], seed: int, offset: int) -> Tuple[Tensor, Tensor]:
    g_x, g_cb = RF.quantize_backward(g_emb, g_res, g_loss, residuals, ids, codebooks, mode, beta, temperature,
                                     gumbel_uniform, seed, offset)
    return g_x, g_cb


@quantize_backward.register_fake
def _(g_emb, g_res, g_loss, residuals, ids, codebooks, mode, beta, temperature, gumbel_uniform, seed, offset):
    Lv, B, D = residuals.shape
    return (torch.empty((B, D), device=residuals.device, dtype=torch.float32), torch.empty_like(codebooks))


def _fwd_setup_context(ctx, inputs, output):
    x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed, offset, impl = inputs
    _emb, res, ids, _loss = output
    ctx.save_for_backward(res, ids, codebooks, gumbel_uniform)
    ctx.cfg = (mode, beta, temperature, seed, offset)
    # a gradient nobody produced (normally the residuals') reaches the kernel as NULL, not as zeros to stream
    ctx.set_materialize_grads(False)


def _fwd_backward(ctx, g_emb, g_res, _g_ids, g_loss):
    res, ids, codebooks, gu = ctx.saved_tensors
    mode, beta, temperature, seed, offset = ctx.cfg
    g_x, g_cb = quantize_backward(g_emb, g_res, g_loss, res, ids, codebooks, mode, beta, temperature, gu, seed, offset)
    return g_x, g_cb, None, None, None, None, None, None, None, None


quantize_forward.register_autograd(_fwd_backward, setup_context=_fwd_setup_context)


# ---- integer side paths used inside RqVae.forward -------------------------------------------------
@torch.library.custom_op(f"{_NS}::corpus_table", mutates_args=(), device_types="cuda")
def corpus_table(sem_ids: Tensor, codebook_size: int) -> Tensor:
    return RF.corpus_table(sem_ids, codebook_size)


@corpus_table.register_fake
def _(sem_ids, codebook_size):
    N, Lv = sem_ids.shape
    return torch.empty((N, Lv + 1), device=sem_ids.device, dtype=torch.int64)


@torch.library.custom_op(f"{_NS}::gather_tokens", mutates_args=(), device_types="cuda")
def gather_tokens(table: Tensor, item_ids: Tensor, seq_mask: Optional[Tensor]) -> Tensor:
    return RF.gather_tokens(table, item_ids, seq_mask)


@gather_tokens.register_fake
def _(table, item_ids, seq_mask):
    W = table.shape[1]
    return torch.empty((*item_ids.shape[:-1], item_ids.shape[-1] * W), device=table.device, dtype=torch.int64)


@torch.library.custom_op(f"{_NS}::kmeans_assign_accumulate", mutates_args=(), device_types="cuda")
def kmeans_assign_accumulate(x: Tensor, centroids: Tensor, impl: int) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
    assign, sums, counts, n_changed = RF.kmeans_assign_accumulate(x, centroids, None, impl)
    return assign, sums, counts, n_changed


@kmeans_assign_accumulate.register_fake
def _(x, centroids, impl):
    B, D = x.shape
    K = centroids.shape[0]
    i64 = dict(device=x.device, dtype=torch.int64)
    return (torch.empty((B,), **i64), torch.empty((K, D), device=x.device, dtype=torch.float32),
            torch.empty((K,), **i64), torch.empty((1,), **i64))


OPS = ("quantize_forward", "quantize_backward", "corpus_table", "gather_tokens", "kmeans_assign_accumulate")
