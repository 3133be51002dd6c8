"""Row sharding of the item corpus across ranks: one process per GPU, codebooks and
encoder replicated, items split into contiguous blocks.  Tokenisation needs no data-path
collective; k-means and the codebook gradient use one all-reduce (torch.distributed / NCCL)."""
from typing import Tuple

import torch


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of `n` rows owned by `rank`; the first n % world ranks get one extra."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_codebook_grads(model, group=None, average: bool = True) -> None:
    """All-reduce of the [L,K,D] codebook gradients (98,304 B at 3x256x32): what DDP does
    implicitly in the reference (train_rqvae.py:115-117,153), as ONE flat NCCL call."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    grads = [l.embedding.weight.grad for l in model.layers if l.embedding.weight.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, group=group)
    if average:
        flat /= dist.get_world_size(group)
    o = 0
    for g in grads:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()
