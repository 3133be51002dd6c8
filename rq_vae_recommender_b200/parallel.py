"""Row sharding of the item corpus across ranks: one process per GPU, codebooks and
encoder replicated, items split into contiguous blocks.  Tokenisation needs no data-path
collective; k-means and the codebook gradient use one all-reduce (torch.distributed / NCCL)."""
from typing import Tuple

import torch


def reserve_sms_for_collectives(n_sms: int = 8) -> None:
    """Leave `n_sms` SMs out of the engine's persistent grids so that an NCCL kernel on another stream
    (codebook-gradient / k-means all-reduce) and the quantiser kernels never wait for each other's SM
    resources.  Must be called before the first kernel launch of the process (the library reads
    RQ_SM_RESERVE once)."""
    import os
    os.environ["RQ_SM_RESERVE"] = str(int(n_sms))


def shard_range(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of `n` rows owned by `rank`; the first n % world ranks get one extra."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_codebook_grads(model, group=None, average: bool = True) -> None:
    """All-reduce of the [L,K,D] codebook gradients (98,304 B at 3x256x32): what DDP does
    implicitly in the reference (train_rqvae.py:115-117,153), as ONE flat NCCL call (three device
    ops in total: pack, all-reduce with in-collective averaging, unpack)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    grads = [l.embedding.weight.grad for l in model.layers if l.embedding.weight.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    nccl = dist.get_backend(group) == "nccl"
    if average and nccl:
        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)
    else:
        dist.all_reduce(flat, group=group)
        if average:
            flat /= dist.get_world_size(group)
    torch._foreach_copy_(grads, [c.view_as(g) for g, c in zip(grads, flat.split([g.numel() for g in grads]))])


def is_distributed(group=None) -> bool:
    import torch.distributed as dist
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


def gather_row_shards(local: torch.Tensor, n: int, group=None) -> torch.Tensor:
    """All-gather of contiguous row shards laid out by `shard_range`: every rank passes its [n_local, ...]
    block and receives the full [n, ...] tensor.  Shards differ by at most one row, so they are padded to
    ceil(n / world) rows for one equal-size all_gather and the padding is dropped afterwards.  The only
    collective of corpus tokenisation: [N/G, L] ids in, [N, L] out."""
    import torch.distributed as dist
    if not is_distributed(group):
        return local
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_range(n, rank, world)
    assert local.shape[0] == hi - lo, f"rank {rank} holds {local.shape[0]} rows, its shard has {hi - lo}"
    per = (n + world - 1) // world
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:hi - lo] = local
    out = torch.empty((world * per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    parts = []
    for r in range(world):
        a, b = shard_range(n, r, world)
        parts.append(out[r * per:r * per + (b - a)])
    return torch.cat(parts, dim=0)
