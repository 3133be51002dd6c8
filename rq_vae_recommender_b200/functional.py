"""Tensor-level wrappers over the C ABI (torch tensors in, torch tensors out).

These are the calls the module mirrors in `modules/`, `init/` make; they add argument checks,
output allocation (torch caching allocator, current stream) and the autograd binding.
"""
from __future__ import annotations

from typing import Optional, Tuple

import os

import torch

from . import _lib as L
from ._lib import (DIST_COSINE, DIST_L2, IMPL_AUTO, IMPL_SIMT, IMPL_TCGEN05, MODE_EVAL,  # noqa: F401
                   MODE_GUMBEL_SOFTMAX, MODE_ROTATION_TRICK, MODE_STE)


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    """CUDA fp32, contiguous, 32-byte aligned (the kernels use 128/256-bit accesses and TMA bulk copies): a view with
    an odd storage offset is cloned instead of being handed to the library misaligned."""
    L.require_cuda(t, name, torch.float32)
    if not t.is_contiguous():
        t = t.contiguous()
    if t.data_ptr() % 32 != 0 and t.numel():
        t = t.clone(memory_format=torch.contiguous_format)
    return t


def quantize_forward(x: torch.Tensor, codebooks: torch.Tensor, mode: int = MODE_EVAL, beta: float = 0.25,
                     temperature: float = 0.2, distance: int = DIST_L2,
                     gumbel_uniform: Optional[torch.Tensor] = None, seed: int = 0, offset: int = 0,
                     want_embeddings: bool = True, want_residuals: bool = True, want_loss: bool = True,
                     impl: int = IMPL_AUTO, gumbel_weights_out: Optional[torch.Tensor] = None):
    """Fused L-level residual quantisation (rq_quantize_forward).

    x [B,D] fp32, codebooks [L,K,D] fp32 (effective).  Returns (emb [L,B,D]|None,
    residuals [L,B,D]|None, ids [L,B] int64, loss [B]|None).  `gumbel_weights_out` ([L, B, Kp] fp32 from
    `gumbel_weights_buffer`) receives the softmax weights of a Gumbel-softmax launch on the tensor cores, for
    `quantize_backward(..., gumbel_weights=)`."""
    x = _f32c(x, "x")
    codebooks = _f32c(codebooks, "codebooks")
    if x.dim() != 2 or codebooks.dim() != 3:
        raise L.RqEngineError("x must be [B,D] and codebooks [L,K,D]")
    B, D = x.shape
    Lv, K, Dc = codebooks.shape
    assert Dc == D, f"codebook dim {Dc} != latent dim {D}"      # modules/quantize.py:100
    if gumbel_uniform is not None:
        gumbel_uniform = _f32c(gumbel_uniform, "gumbel_uniform")
        if tuple(gumbel_uniform.shape) != (Lv, B, K):
            raise L.RqEngineError(f"gumbel_uniform must be [L,B,K]={Lv, B, K}, got {tuple(gumbel_uniform.shape)}")
    dev = x.device
    with torch.cuda.device(dev):
        emb = torch.empty((Lv, B, D), device=dev, dtype=torch.float32) if want_embeddings else None
        res = torch.empty((Lv, B, D), device=dev, dtype=torch.float32) if want_residuals else None
        ids = torch.empty((Lv, B), device=dev, dtype=torch.int64)
        loss = torch.empty((B,), device=dev, dtype=torch.float32) if want_loss else None
        rc = L.lib().rq_quantize_forward_ws(L.ptr(x), L.ptr(codebooks), B, D, K, Lv, int(mode), int(distance),
                                            float(beta), float(temperature), L.ptr(gumbel_uniform), int(seed),
                                            int(offset), L.ptr(emb), L.ptr(res), L.ptr(ids), L.ptr(loss),
                                            L.ptr(gumbel_weights_out), int(impl), L.stream_of(x))
    L.check(rc, "rq_quantize_forward")
    return emb, res, ids, loss


def gumbel_weights_buffer(x: torch.Tensor, codebooks: torch.Tensor, mode: int, distance: int = DIST_L2,
                          impl: int = IMPL_AUTO) -> Optional[torch.Tensor]:
    """The [L, B, Kp] buffer `quantize_forward(..., gumbel_weights_out=)` fills when the launch takes the tensor-core
    Gumbel-softmax path (rq_gumbel_weights_saved), else None.  What autograd keeps alive in the reference: the output of
    F.softmax (modules/quantize.py:127), 4 B K bytes per level."""
    B, D = x.shape
    Lv, K, _ = codebooks.shape
    if not L.lib().rq_gumbel_weights_saved(B, D, K, Lv, int(mode), int(distance), int(impl)):
        return None
    return torch.empty((Lv, B, (K + 3) & ~3), device=x.device, dtype=torch.float32)


def quantize_backward(g_emb, g_res, g_loss, residuals, ids, codebooks, mode, beta, temperature,
                      gumbel_uniform=None, seed: int = 0, offset: int = 0,
                      g_codebooks_out: Optional[torch.Tensor] = None,
                      accumulate: bool = False, thread_per_row: bool = False,
                      gumbel_weights: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """rq_quantize_backward: returns (g_x [B,D], g_codebooks [L,K,D]).  `g_codebooks_out` (contiguous fp32
    [L,K,D]) receives the codebook gradient in place — e.g. a persistent all-reduce bucket.  `accumulate=True` adds
    to what `g_codebooks_out` holds instead of overwriting it (gradient accumulation; with a buffer zeroed BEFORE the
    forward no memset sits between the forward and the backward kernel, so the two overlap through programmatic
    dependent launch)."""
    residuals = _f32c(residuals, "residuals")
    codebooks = _f32c(codebooks, "codebooks")
    L.require_cuda(ids, "ids", torch.int64)
    ids = ids.contiguous()
    Lv, B, D = residuals.shape
    K = codebooks.shape[1]
    g_emb = None if g_emb is None else _f32c(g_emb, "g_emb")
    g_res = None if g_res is None else _f32c(g_res, "g_res")
    g_loss = None if g_loss is None else _f32c(g_loss, "g_loss")
    gumbel_uniform = None if gumbel_uniform is None else _f32c(gumbel_uniform, "gumbel_uniform")
    dev = residuals.device
    with torch.cuda.device(dev):
        g_x = torch.empty((B, D), device=dev, dtype=torch.float32)
        if g_codebooks_out is None:
            g_cb = torch.empty((Lv, K, D), device=dev, dtype=torch.float32)
        else:
            g_cb = L.require_cuda(g_codebooks_out, "g_codebooks_out", torch.float32)
            if tuple(g_cb.shape) != (Lv, K, D) or not g_cb.is_contiguous():
                raise L.RqEngineError(f"g_codebooks_out must be contiguous [L,K,D]={Lv, K, D}")
        if accumulate and g_codebooks_out is None:
            raise L.RqEngineError("quantize_backward: accumulate=True needs g_codebooks_out")
        if gumbel_weights is not None:
            gumbel_weights = L.require_cuda(gumbel_weights, "gumbel_weights", torch.float32)
            if tuple(gumbel_weights.shape) != (Lv, B, (K + 3) & ~3) or not gumbel_weights.is_contiguous():
                raise L.RqEngineError(f"gumbel_weights must be contiguous [L,B,Kp]={Lv, B, (K + 3) & ~3}")
        rc = L.lib().rq_quantize_backward_ws(L.ptr(g_emb), L.ptr(g_res), L.ptr(g_loss), L.ptr(residuals), L.ptr(ids),
                                             L.ptr(codebooks), B, D, K, Lv, int(mode), float(beta), float(temperature),
                                             L.ptr(gumbel_uniform), int(seed), int(offset), L.ptr(g_x), L.ptr(g_cb),
                                             L.ptr(gumbel_weights),
                                             (1 if accumulate else 0) | (2 if thread_per_row else 0),
                                             L.stream_of(residuals))
    L.check(rc, "rq_quantize_backward")
    return g_x, g_cb


class ResidualQuantizeFn(torch.autograd.Function):
    """Autograd binding of the fused forward/backward pair.  Stands in for the ~45-node per-level
    autograd graph the reference builds implicitly."""

    @staticmethod
    def forward(ctx, x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed, offset, impl):
        # the codebook gradient is zeroed HERE, ahead of the forward kernel: nothing but the caller's own kernels then
        # sits between the fused forward and the fused backward, whose launch PDL overlaps with its predecessor's tail
        ctx.g_cb = torch.zeros_like(codebooks) if mode != MODE_GUMBEL_SOFTMAX else None
        # Gumbel-softmax on the tensor cores: the softmax weights are kept for the backward, as autograd keeps the output
        # of F.softmax in the reference (RQ_GUMBEL_RECOMPUTE=1 trades them for a recomputation: 4 B K bytes per level)
        w_buf = None
        if mode == MODE_GUMBEL_SOFTMAX and not os.environ.get("RQ_GUMBEL_RECOMPUTE"):
            w_buf = gumbel_weights_buffer(x, codebooks, mode, distance, impl)
        emb, res, ids, loss = quantize_forward(x, codebooks, mode, beta, temperature, distance, gumbel_uniform,
                                               seed, offset, True, True, True, impl, gumbel_weights_out=w_buf)
        ctx.gumbel_w = w_buf
        ctx.save_for_backward(res, ids, codebooks, gumbel_uniform if gumbel_uniform is not None else x.new_empty(0))
        ctx.cfg = (mode, beta, temperature, gumbel_uniform is not None, seed, offset)
        ctx.mark_non_differentiable(ids)
        # an output nobody differentiated (normally the residuals) reaches backward as None, not as a
        # zero tensor the kernel would have to stream
        ctx.set_materialize_grads(False)
        return emb, res, ids, loss

    @staticmethod
    def backward(ctx, g_emb, g_res, _g_ids, g_loss):
        res, ids, codebooks, gu = ctx.saved_tensors
        mode, beta, temperature, has_gu, seed, offset = ctx.cfg
        g_cb_buf, ctx.g_cb = ctx.g_cb, None
        w_buf, ctx.gumbel_w = ctx.gumbel_w, None
        g_x, g_cb = quantize_backward(g_emb, g_res, g_loss, res, ids, codebooks, mode, beta, temperature,
                                      gu if has_gu else None, seed, offset, g_codebooks_out=g_cb_buf,
                                      accumulate=g_cb_buf is not None, gumbel_weights=w_buf)
        return g_x, g_cb, None, None, None, None, None, None, None, None


def residual_quantize(x, codebooks, mode=MODE_EVAL, beta=0.25, temperature=0.2, distance=DIST_L2,
                      gumbel_uniform=None, seed=0, offset=0, impl=IMPL_AUTO):
    """Differentiable fused quantiser: (emb [L,B,D], residuals [L,B,D], ids [L,B], loss [B]).

    The engine computes in fp32 ("all arithmetic is fp32"): half / bfloat16 latents (the reference runs its
    encoder under `accelerator.autocast()`, train_rqvae.py:147) are upcast here, outputs and gradients are fp32 like
    the reference's quantiser, whose distance GEMM autocast keeps in fp32 only by accident of `torch.cdist`-free code."""
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
        x = x.float()
    if codebooks.is_cuda and codebooks.dtype in (torch.float16, torch.bfloat16):
        codebooks = codebooks.float()
    if (mode == MODE_GUMBEL_SOFTMAX and distance == DIST_COSINE and torch.is_grad_enabled()
            and (x.requires_grad or codebooks.requires_grad)):
        raise L.RqEngineError(
            "GUMBEL_SOFTMAX with COSINE distance is forward-only in the engine: the fused backward differentiates the "
            "L2 logits (quantize.py:108-112).  Use distance_mode=L2 (what RqVae always does, rqvae.py:65-73), or call "
            "under torch.no_grad()")
    if torch.compiler.is_compiling():
        # inside torch.compile the ctypes call is untraceable: go through the registered custom op (ops.py), which
        # carries its own fake implementation and autograd formula
        return _ops.quantize_forward(x, codebooks, int(mode), float(beta), float(temperature), int(distance),
                                     gumbel_uniform, int(seed), int(offset), int(impl))
    if torch.is_grad_enabled() and (x.requires_grad or codebooks.requires_grad):
        return ResidualQuantizeFn.apply(x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed,
                                        offset, impl)
    return quantize_forward(x, codebooks, mode, beta, temperature, distance, gumbel_uniform, seed, offset,
                            True, True, True, impl)


# ---- k-means -------------------------------------------------------------------------------
def kmeans_assign_accumulate(x, centroids, assign: Optional[torch.Tensor] = None, impl: int = IMPL_AUTO):
    """One assignment+accumulation pass.  Returns (assign [B] i64, sums [K,D], counts [K] i64,
    n_changed [1] i64).  `assign`, when given, is the previous assignment and is updated in place."""
    x = _f32c(x, "x")
    centroids = _f32c(centroids, "centroids")
    B, D = x.shape
    K = centroids.shape[0]
    dev = x.device
    with torch.cuda.device(dev):
        has_prev = assign is not None
        if not has_prev:
            assign = torch.empty((B,), device=dev, dtype=torch.int64)
        else:
            L.require_cuda(assign, "assign", torch.int64)
        sums = torch.empty((K, D), device=dev, dtype=torch.float32)
        counts = torch.empty((K,), device=dev, dtype=torch.int64)
        n_changed = torch.empty((1,), device=dev, dtype=torch.int64)
        rc = L.lib().rq_kmeans_assign_accumulate(L.ptr(x), L.ptr(centroids), B, D, K, L.ptr(assign), int(has_prev),
                                                 L.ptr(sums), L.ptr(counts), L.ptr(n_changed), int(impl),
                                                 L.stream_of(x))
    L.check(rc, "rq_kmeans_assign_accumulate")
    return assign, sums, counts, n_changed


def kmeans_update(sums, counts, centroids):
    """In-place centroid update.  Returns (max_shift [1] f32, n_empty [1] i64, empty_ids [K] i64)."""
    K, D = centroids.shape
    dev = centroids.device
    with torch.cuda.device(dev):
        max_shift = torch.empty((1,), device=dev, dtype=torch.float32)
        n_empty = torch.empty((1,), device=dev, dtype=torch.int64)
        empty_ids = torch.empty((K,), device=dev, dtype=torch.int64)
        rc = L.lib().rq_kmeans_update(L.ptr(sums), L.ptr(counts), L.ptr(centroids), D, K, L.ptr(max_shift),
                                      L.ptr(n_empty), L.ptr(empty_ids), L.stream_of(centroids))
    L.check(rc, "rq_kmeans_update")
    return max_shift, n_empty, empty_ids


# ---- tokenizer integer paths ------------------------------------------------------------------
def corpus_table(sem_ids: torch.Tensor, codebook_size: int) -> torch.Tensor:
    """[N,L] int64 (any strides) -> [N,L+1] int64 = [ids | dedup] (semids.py:86-99)."""
    if torch.compiler.is_compiling():
        return _ops.corpus_table(sem_ids, int(codebook_size))
    L.require_cuda(sem_ids, "sem_ids", torch.int64)
    N, Lv = sem_ids.shape
    dev = sem_ids.device
    with torch.cuda.device(dev):
        table = torch.empty((N, Lv + 1), device=dev, dtype=torch.int64)
        nbytes = int(L.lib().rq_corpus_table_workspace_bytes(N, Lv))
        ws = torch.empty((nbytes,), device=dev, dtype=torch.uint8)
        rc = L.lib().rq_corpus_table(L.ptr(sem_ids), sem_ids.stride(0), sem_ids.stride(1), N, Lv,
                                     int(codebook_size), L.ptr(table), L.ptr(ws), nbytes, L.stream_of(sem_ids))
    L.check(rc, "rq_corpus_table")
    return table


def id_diversity(table: torch.Tensor, codebook_size: int):
    """[N,L+1] int64 corpus table -> (entropy [] float64, used [L] int64, max_dup [] int64), device tensors
    (rq_id_diversity; train_rqvae.py:227-237)."""
    L.require_cuda(table, "table", torch.int64)
    table = table.contiguous()
    N, W = table.shape
    if N < 1 or W < 2:
        raise L.RqEngineError("id_diversity: table must be [N >= 1, L + 1]")
    dev = table.device
    with torch.cuda.device(dev):
        entropy = torch.empty((), device=dev, dtype=torch.float64)
        used = torch.empty((W - 1,), device=dev, dtype=torch.int64)
        max_dup = torch.empty((), device=dev, dtype=torch.int64)
        nbytes = int(L.lib().rq_id_diversity_workspace_bytes(W - 1, int(codebook_size)))
        ws = torch.empty((max(nbytes, 4),), device=dev, dtype=torch.uint8)
        rc = L.lib().rq_id_diversity(L.ptr(table), N, W - 1, int(codebook_size), L.ptr(entropy), L.ptr(used),
                                     L.ptr(max_dup), L.ptr(ws), nbytes, L.stream_of(table))
    L.check(rc, "rq_id_diversity")
    return entropy, used, max_dup


def prefix_index(table: torch.Tensor, p: int, radix: int) -> torch.Tensor:
    """Sorted packed keys of table[:, :p] (uint64 stored in an int64 tensor)."""
    L.require_cuda(table, "table", torch.int64)
    table = table.contiguous()
    N, W = table.shape
    dev = table.device
    with torch.cuda.device(dev):
        keys = torch.empty((N,), device=dev, dtype=torch.int64)
        nbytes = int(L.lib().rq_prefix_index_workspace_bytes(N))
        ws = torch.empty((nbytes,), device=dev, dtype=torch.uint8)
        rc = L.lib().rq_prefix_index_build(L.ptr(table), N, W, int(p), int(radix), L.ptr(keys), L.ptr(ws), nbytes,
                                           L.stream_of(table))
    L.check(rc, "rq_prefix_index_build")
    return keys


def prefix_exists(sorted_keys: torch.Tensor, queries: torch.Tensor, radix: int) -> torch.Tensor:
    """queries [..., p] int64 -> bool [...]."""
    L.require_cuda(queries, "queries", torch.int64)
    p = queries.shape[-1]
    q = queries.reshape(-1, p).contiguous()
    dev = queries.device
    with torch.cuda.device(dev):
        out = torch.empty((q.shape[0],), device=dev, dtype=torch.uint8)
        rc = L.lib().rq_prefix_exists(L.ptr(sorted_keys), sorted_keys.shape[0], L.ptr(q), q.shape[0], p, int(radix),
                                      L.ptr(out), L.stream_of(queries))
    L.check(rc, "rq_prefix_exists")
    return out.view(torch.bool).reshape(queries.shape[:-1])


def gather_tokens(table: torch.Tensor, item_ids: torch.Tensor, seq_mask: Optional[torch.Tensor] = None):
    """table [N,W], item_ids [B,n] -> [B, n*W] with padding -> -1 (semids.py:122-123, 137-141)."""
    if torch.compiler.is_compiling():
        return _ops.gather_tokens(table, item_ids, seq_mask)
    L.require_cuda(table, "table", torch.int64)
    L.require_cuda(item_ids, "item_ids", torch.int64)
    table = table.contiguous()
    ids = item_ids.contiguous()
    Bn, n = ids.shape
    W = table.shape[1]
    m = None
    if seq_mask is not None:
        m = seq_mask.to(torch.bool).contiguous().view(torch.uint8)
    dev = table.device
    with torch.cuda.device(dev):
        out = torch.empty((Bn, n * W), device=dev, dtype=torch.int64)
        rc = L.lib().rq_gather_tokens(L.ptr(table), table.shape[0], W, L.ptr(ids), L.ptr(m), Bn * n, L.ptr(out),
                                      L.stream_of(table))
    L.check(rc, "rq_gather_tokens")
    return out


from . import ops as _ops  # noqa: E402  (torch.library registration; imports this module back)


ACT_NONE, ACT_SILU, ACT_L2NORM, ACT_TF32_OUT, ACT_CTA_PAIR = 0, 1, 2, 0x10, 0x20


def encoder_layer_supported(rows: int, in_features: int, out_features: int, act: int) -> bool:
    return bool(L.lib().rq_encoder_layer_supported(int(rows), int(in_features), int(out_features), int(act)))


def encoder_layer(x: torch.Tensor, weight: torch.Tensor, act: int = ACT_NONE, out: Optional[torch.Tensor] = None):
    """act(x @ weight.T) for one bias-free Linear of the MLP (rq_encoder_layer; modules/encoder.py:17-32): TF32 tensor-core
    product with SiLU / the final L2 normalisation fused.  x [rows, in], weight [out, in] fp32 CUDA; no autograd."""
    x = _f32c(x, "x")
    weight = _f32c(weight, "weight")
    if x.dim() != 2 or weight.dim() != 2 or x.shape[1] != weight.shape[1]:
        raise L.RqEngineError(f"encoder_layer: x {tuple(x.shape)} and weight {tuple(weight.shape)} do not match")
    rows, n_in = x.shape
    n_out = weight.shape[0]
    if out is None:
        out = torch.empty((rows, n_out), device=x.device, dtype=torch.float32)
    elif out.shape != (rows, n_out) or out.dtype != torch.float32 or not out.is_contiguous() or out.device != x.device:
        raise L.RqEngineError("encoder_layer: out must be a contiguous fp32 [rows, out_features] tensor on x's device")
    with torch.cuda.device(x.device):
        rc = L.lib().rq_encoder_layer(L.ptr(x), L.ptr(weight), L.ptr(out), rows, n_in, n_out, int(act), L.stream_of(x))
    L.check(rc, "rq_encoder_layer")
    return out
