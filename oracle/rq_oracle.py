"""CPU ORACLE for the residual-quantization hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, op for op, what AdamLTy/RQ-VAE-Recommender computes on the path
BASELINE.json's `north_star` names.  It is the checker: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference` legs may
import it.  The product (`rq_vae_recommender_b200`) never does and has no CPU fallback.

Arithmetic is fp32 through torch's CPU ATen ops (the same MKL SGEMM / reductions the
reference dispatches to on CPU), integer work is numpy.  Parity status: PINNED — the
reference has no golden vectors of its own, so `oracle/make_golden.py`
imports the unmodified reference in the build container, runs it on seeded inputs and
commits inputs+outputs under `tests/golden/`; `tests/test_oracle_golden.py` checks this
file against those fixtures bit for bit (ids, dedup column) / to 0 ulp (floats).

Every function cites the reference file:line it follows (paths relative to the
reference root).
"""
from __future__ import annotations

import math
from typing import List, NamedTuple, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

# --- a1: enums (modules/quantize.py:16-25) ---------------------------------------------
MODE_GUMBEL_SOFTMAX = 1
MODE_STE = 2
MODE_ROTATION_TRICK = 3
MODE_EVAL = 0  # `not self.training` branch, modules/quantize.py:148-150
DIST_L2 = 1
DIST_COSINE = 2


class LevelOut(NamedTuple):
    emb_out: torch.Tensor  # [B, D]   value forwarded to the next level / decoder
    ids: torch.Tensor      # [B]      int64
    loss: torch.Tensor     # [B]
    emb: torch.Tensor      # [B, D]   "value" that enters QuantizeLoss
    dist: torch.Tensor     # [B, K]


class RqOut(NamedTuple):
    embeddings: torch.Tensor  # [L, B, D]  (reference returns the [B, D, L] view of this memory)
    residuals: torch.Tensor   # [L, B, D]
    sem_ids: torch.Tensor     # [L, B] int64 (reference returns the [B, L] view)
    quantize_loss: torch.Tensor  # [B]


# --- a3: distances (modules/quantize.py:107-119) -------------------------------------------
def l2_distances(x: torch.Tensor, codebook: torch.Tensor) -> torch.Tensor:
    """modules/quantize.py:108-112 — ||x||^2 + ||c||^2 - (2x)@C^T, in that association."""
    return (
        (x ** 2).sum(axis=1, keepdim=True)
        + (codebook.T ** 2).sum(axis=0, keepdim=True)
        - 2 * x @ codebook.T
    )


def cosine_distances(x: torch.Tensor, codebook: torch.Tensor) -> torch.Tensor:
    """modules/quantize.py:114-117 — note the reference's association:
    ((x/||x||) @ C^T) / ||c||, negated."""
    return -(
        x / x.norm(dim=1, keepdim=True) @ (codebook.T) / codebook.T.norm(dim=0, keepdim=True)
    )


def sample_gumbel_from_uniform(u: torch.Tensor, eps: float = 1e-20) -> torch.Tensor:
    """distributions/gumbel.py:7-10 with the uniform draw `U` made explicit."""
    return -torch.log(-torch.log(u + eps) + eps)


def gumbel_softmax_weights(logits: torch.Tensor, temperature: float, u: torch.Tensor) -> torch.Tensor:
    """distributions/gumbel.py:13-17."""
    y = logits + sample_gumbel_from_uniform(u)
    return F.softmax(y / temperature, dim=-1)


def rotation_trick(x: torch.Tensor, emb: torch.Tensor) -> torch.Tensor:
    """modules/quantize.py:34-45 and :133-142.  `x` carries grad; u, q, w, lambda detached."""
    u = x / (x.norm(dim=-1, keepdim=True) + 1e-8)
    q = emb / (emb.norm(dim=-1, keepdim=True) + 1e-8)
    e = x.unsqueeze(1)                                       # b 1 d
    w = F.normalize(u + q, p=2, dim=1, eps=1e-6).detach()
    rot = (
        e
        - 2 * (e @ w.unsqueeze(2) @ w.unsqueeze(1))
        + 2 * (e @ u.unsqueeze(2).detach() @ q.unsqueeze(1).detach())
    ).squeeze()
    lam = (torch.norm(emb, dim=1, keepdim=True) / (torch.norm(x, dim=1, keepdim=True) + 1e-6)).detach()
    return rot * lam


def quantize_loss(query: torch.Tensor, value: torch.Tensor, beta: float) -> torch.Tensor:
    """modules/loss.py:39-42."""
    emb_loss = ((query.detach() - value) ** 2).sum(axis=[-1])
    query_loss = ((query - value.detach()) ** 2).sum(axis=[-1])
    return emb_loss + beta * query_loss


def quantize_level(
    x: torch.Tensor,
    codebook: torch.Tensor,
    mode: int = MODE_EVAL,
    beta: float = 0.25,
    temperature: float = 0.2,
    distance: int = DIST_L2,
    uniform: Optional[torch.Tensor] = None,
) -> LevelOut:
    """One `Quantize.forward` (modules/quantize.py:99-156) on an *effective* codebook
    (`out_proj(embedding.weight)`, :105).  `mode == MODE_EVAL` is the `not self.training`
    branch (:148-150); the three training estimators are :124-142."""
    assert x.shape[-1] == codebook.shape[-1]                 # :100
    if distance == DIST_L2:
        dist = l2_distances(x, codebook)
    elif distance == DIST_COSINE:
        dist = cosine_distances(x, codebook)
    else:
        raise Exception("Unsupported Quantize distance mode.")   # :119
    _, ids = (dist.detach()).min(axis=1)                     # :121 (first index wins ties)
    if mode == MODE_EVAL:
        emb_out = codebook[ids]                              # :149
        emb = emb_out
    elif mode == MODE_GUMBEL_SOFTMAX:
        assert uniform is not None, "oracle needs the uniform draw made explicit"
        weights = gumbel_softmax_weights(-dist, temperature, uniform)   # :125-127
        emb = weights @ codebook                             # :128
        emb_out = emb
    elif mode == MODE_STE:
        emb = codebook[ids]                                  # :131
        emb_out = x + (emb - x).detach()                     # :132
    elif mode == MODE_ROTATION_TRICK:
        emb = codebook[ids]                                  # :134
        emb_out = rotation_trick(x, emb)                     # :135-142
    else:
        raise Exception("Unsupported Quantize forward mode.")    # :144
    loss = quantize_loss(x, emb, beta)                       # :146 / :150
    return LevelOut(emb_out, ids, loss, emb, dist)


def rq_forward(
    latents: torch.Tensor,
    codebooks: Sequence[torch.Tensor],
    mode: int = MODE_EVAL,
    beta: float = 0.25,
    temperature: float = 0.2,
    distance: int = DIST_L2,
    uniforms: Optional[Sequence[torch.Tensor]] = None,
) -> RqOut:
    """The level loop of `RqVae.get_semantic_ids` (modules/rqvae.py:119-138) after the
    encoder: residuals.append(res); quantize; loss +=; res = res - emb."""
    res = latents
    qloss = 0
    embs: List[torch.Tensor] = []
    residuals: List[torch.Tensor] = []
    sem_ids: List[torch.Tensor] = []
    for l, cb in enumerate(codebooks):
        residuals.append(res)
        out = quantize_level(res, cb, mode, beta, temperature, distance,
                             None if uniforms is None else uniforms[l])
        qloss = qloss + out.loss
        res = res - out.emb_out
        sem_ids.append(out.ids)
        embs.append(out.emb_out)
    return RqOut(torch.stack(embs), torch.stack(residuals), torch.stack(sem_ids), qloss)


def rq_forward_backward(latents, codebooks, mode, beta, temperature, g_emb, g_loss,
                        uniforms=None, g_res=None):
    """Autograd through `rq_forward` — what the reference's implicit autograd graph
     produces for upstream gradients on embeddings [L,B,D],
    residuals [L,B,D] (optional) and quantize_loss [B]."""
    x = latents.detach().clone().requires_grad_(True)
    cbs = [c.detach().clone().requires_grad_(True) for c in codebooks]
    out = rq_forward(x, cbs, mode, beta, temperature, DIST_L2, uniforms)
    obj = (out.embeddings * g_emb).sum() + (out.quantize_loss * g_loss).sum()
    if g_res is not None:
        obj = obj + (out.residuals * g_res).sum()
    obj.backward()
    gcb = torch.stack([c.grad if c.grad is not None else torch.zeros_like(c) for c in cbs])
    return out, x.grad, gcb


# --- tie band (north_star: "bit-exact except where the fp32 top-2 gap is under 1e-6 rel") ---
def tie_band(x: torch.Tensor, codebook: torch.Tensor, rel: float = 1e-6):
    """Returns (in_band_def1, in_band_def2, gap, d1): per row, whether the fp32 top-2
    distance gap is < rel*d1 (definition 1) or < rel*(||x||^2+||c||^2) (definition 2,
    the absolute error scale of the expanded form).  Computed in
    float64 from the direct form so it is independent of the implementation under test."""
    x64, c64 = x.double(), codebook.double()
    d = torch.cdist(x64, c64, p=2) ** 2
    top2 = torch.topk(d, 2, dim=1, largest=False)
    d1, d2 = top2.values[:, 0], top2.values[:, 1]
    gap = d2 - d1
    scale2 = (x64 ** 2).sum(1) + (c64[top2.indices[:, 0]] ** 2).sum(1)
    return gap < rel * d1, gap < rel * scale2, gap, d1


# --- a12: k-means (init/kmeans.py) ------------------------------------------------------------
def kmeans_assign(x: torch.Tensor, centroids: torch.Tensor, chunk: int = 4096) -> torch.Tensor:
    """init/kmeans.py:41-44 — direct form sum_d (x-c)^2, argmin over clusters.  Chunked
    over rows so the [B,K,D] temporary stays small; per-row arithmetic is unchanged."""
    out = []
    for s in range(0, x.shape[0], chunk):
        xb = x[s:s + chunk]
        d = ((xb.unsqueeze(1) - centroids.unsqueeze(0)) ** 2).sum(axis=2)
        out.append(d.min(axis=1).indices)
    return torch.cat(out) if out else torch.zeros(0, dtype=torch.int64)


def kmeans_update(x: torch.Tensor, centroids: torch.Tensor, reseed_rows=None):
    """One `_update_centroids` (init/kmeans.py:40-58).  Empty cluster -> a random row of x
    (:51-53); the draw is `torch.randint(0, B, (1,))` per empty cluster in cluster order —
    made explicit through `reseed_rows` (a list consumed in order) when given."""
    if x.size(0) == 0:
        raise ValueError("Can not choose random element from x, x is empty")   # :55
    idx = kmeans_assign(x, centroids)
    k = centroids.shape[0]
    new_c = centroids.clone()
    counts = torch.bincount(idx, minlength=k)
    it = iter(reseed_rows) if reseed_rows is not None else None
    for cluster in range(k):
        if counts[cluster] == 0:
            row = int(next(it)) if it is not None else int(torch.randint(0, x.size(0), (1,)))
            new_c[cluster, :] = x[row]
        else:
            new_c[cluster, :] = x[idx == cluster, :].mean(axis=0)    # :57
    return new_c, idx


def kmeans_run(x: torch.Tensor, k: int, init_idx=None, max_iters=None, stop_threshold=1e-10):
    """`Kmeans.run` (init/kmeans.py:60-74).  `init_idx` = the K distinct rows that
    `np.random.choice(B, k, replace=False)` (:36) returned; drawn here if None."""
    if init_idx is None:
        init_idx = np.random.choice(x.shape[0], k, replace=False)
    centroids = x[torch.as_tensor(np.asarray(init_idx))].clone()
    assignment = None
    i = 0
    while max_iters is None or i < max_iters:
        old_c = centroids.clone()
        centroids, assignment = kmeans_update(x, centroids)
        if torch.norm(centroids - old_c, dim=1).max() < stop_threshold:
            break
        i += 1
    return centroids, assignment, i


# --- a14/a15/a13: tokenizer integer paths (modules/tokenizer/semids.py), numpy -------------------
def corpus_dedup_column(sem_ids: np.ndarray) -> np.ndarray:
    """semids.py:74-101 — for item i the dedup value is the number of EARLIER corpus items
    carrying the same L-tuple (in-batch lower-triangle hits :86-87 plus hits against all
    previously cached batches :93-94), i.e. a stable cumulative count per key."""
    sem_ids = np.asarray(sem_ids, dtype=np.int64)
    seen = {}
    out = np.zeros(sem_ids.shape[0], dtype=np.int64)
    for i, row in enumerate(map(tuple, sem_ids.tolist())):
        c = seen.get(row, 0)
        out[i] = c
        seen[row] = c + 1
    return out


def corpus_ids_table(sem_ids: np.ndarray) -> np.ndarray:
    """semids.py:98-99 — cached_ids = [ids | dedup] as int64 [N, L+1]."""
    sem_ids = np.asarray(sem_ids, dtype=np.int64)
    return np.concatenate([sem_ids, corpus_dedup_column(sem_ids)[:, None]], axis=1)


def id_diversity(corpus_ids: torch.Tensor, vae_n_layers: int, vae_codebook_size: int) -> dict:
    """The eval-time id-diversity log, op for op as train_rqvae.py:227-237 computes it (torch.unique over the
    tuple rows and over each column; float32 arithmetic)."""
    corpus_ids = torch.as_tensor(corpus_ids)
    log = {}
    max_duplicates = corpus_ids[:, -1].max() / corpus_ids.shape[0]                      # :228
    _, counts = torch.unique(corpus_ids[:, :-1], dim=0, return_counts=True)             # :230
    p = counts / corpus_ids.shape[0]                                                    # :231
    rqvae_entropy = -(p * torch.log(p)).sum()                                           # :232
    for cid in range(vae_n_layers):                                                     # :234-236
        _, counts = torch.unique(corpus_ids[:, cid], return_counts=True)
        log[f"codebook_usage_{cid}"] = len(counts) / vae_codebook_size
    log["rqvae_entropy"] = rqvae_entropy.cpu().item()                                   # :238
    log["max_id_duplicates"] = max_duplicates.cpu().item()                              # :239
    return log


def exists_prefix(cached_ids: np.ndarray, prefix: np.ndarray, strict_reference_quirk: bool = False) -> np.ndarray:
    """semids.py:103-120.  `strict_reference_quirk=True` reproduces `math.ceil(n // 16)`
    (:114): only full chunks of 16 along dim 0 are evaluated, the remainder stays False."""
    if cached_ids is None:
        raise Exception("No match can be found in empty cache.")     # :107
    prefix = np.asarray(prefix, dtype=np.int64)
    p = prefix.shape[-1]
    keys = {tuple(r) for r in np.asarray(cached_ids)[:, :p].tolist()}
    flat = prefix.reshape(-1, p)
    res = np.fromiter((tuple(r) in keys for r in flat.tolist()), dtype=bool, count=flat.shape[0])
    res = res.reshape(prefix.shape[:-1])
    if strict_reference_quirk:
        n_eval = (prefix.shape[0] // 16) * 16
        res = res.copy()
        res[n_eval:] = False
    return res


def tokenize_from_cache(cached_ids: np.ndarray, ids: np.ndarray, seq_mask: Optional[np.ndarray] = None) -> np.ndarray:
    """semids.py:122-123 and :137-141 — gather rows, flatten to [B, N*(L+1)], pad -> -1."""
    cached_ids = np.asarray(cached_ids)
    ids = np.asarray(ids)
    b, n = ids.shape
    out = cached_ids[ids.reshape(-1)].reshape(b, n * cached_ids.shape[1]).copy()
    if seq_mask is not None:
        m = np.repeat(np.asarray(seq_mask, dtype=bool), cached_ids.shape[1], axis=1)
        out[~m] = -1
    return out


# --- encoder (modules/encoder.py:7-36) for the end-to-end tokenisation oracle ---------------------
def mlp_forward(x: torch.Tensor, weights: Sequence[torch.Tensor], normalize: bool) -> torch.Tensor:
    """Bias-free Linear/SiLU stack, optional trailing L2 normalisation (encoder.py:23-32)."""
    h = x
    for i, w in enumerate(weights):
        h = F.linear(h, w)
        if i != len(weights) - 1:
            h = F.silu(h)
    if normalize:
        h = F.normalize(h, p=2, dim=-1, eps=1e-12)           # modules/normalize.py:7-8
    return h


def effective_codebook(weight: torch.Tensor, sim_vq_weight: Optional[torch.Tensor], normalize: bool) -> torch.Tensor:
    """`out_proj(embedding.weight)` (modules/quantize.py:70-73, :105)."""
    c = weight
    if sim_vq_weight is not None:
        c = F.linear(c, sim_vq_weight)
    if normalize:
        c = F.normalize(c, p=2, dim=-1, eps=1e-12)
    return c
