"""Row sharding of the item corpus across ranks: one process per GPU, codebooks and
encoder replicated, items split into contiguous blocks.  Tokenisation needs no data-path
collective; k-means and the codebook gradient use one all-reduce: torch.distributed / NCCL, or the engine's own
one-shot kernel over peer-mapped memory (`PeerAllReduce`)."""
from typing import Optional, Tuple

import torch


def reserve_sms_for_collectives(n_sms: int = 8) -> None:
    """Leave `n_sms` SMs out of the engine's persistent grids so that an NCCL kernel on another stream
    (codebook-gradient / k-means all-reduce) and the quantiser kernels never wait for each other's SM
    resources.  Must be called before the first kernel launch of the process (the library reads
    RQ_SM_RESERVE once)."""
    import os
    os.environ["RQ_SM_RESERVE"] = str(int(n_sms))


def gpu_numa_node(device=None) -> Optional[int]:
    """NUMA node the GPU hangs off (sysfs: its PCI function's `numa_node`), or None when the platform exposes none."""
    try:
        idx = torch.cuda.current_device() if device is None else torch.device(device).index
        if idx is None:
            idx = torch.cuda.current_device()
        pr = torch.cuda.get_device_properties(idx)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except (OSError, ValueError, AttributeError, RuntimeError, AssertionError):
        return None


class host_memory_near_gpu:
    """Context manager: host memory allocated inside the block by the calling thread (pinned staging buffers:
    `tensor.pin_memory()`, `cudaHostAlloc`) is placed on the GPU's own NUMA node (`set_mempolicy(MPOL_PREFERRED)`),
    so that the per-step H2D / D2H copies of a one-process-per-GPU job do not cross the inter-socket link.  CPU
    affinity is left alone.  `.node` is the node used, None when nothing was changed (no NUMA information, or a
    platform without the syscall)."""
    _SYSCALL = {"x86_64": 238, "aarch64": 237}
    _MPOL_DEFAULT, _MPOL_PREFERRED = 0, 1

    def __init__(self, device=None):
        self.node = gpu_numa_node(device)
        self._libc = None

    def _set(self, mode: int, node: Optional[int]) -> bool:
        import ctypes
        import platform
        nr = self._SYSCALL.get(platform.machine())
        if nr is None:
            return False
        if self._libc is None:
            self._libc = ctypes.CDLL(None, use_errno=True)
        if node is None:
            return self._libc.syscall(nr, mode, None, 0) == 0
        words = node // 64 + 1
        mask = (ctypes.c_ulong * words)()
        mask[node // 64] = 1 << (node % 64)
        return self._libc.syscall(nr, mode, mask, words * 64 + 1) == 0

    def __enter__(self):
        if self.node is not None and not self._set(self._MPOL_PREFERRED, self.node):
            self.node = None
        return self

    def __exit__(self, *exc):
        if self.node is not None:
            self._set(self._MPOL_DEFAULT, None)
        return False


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


class PeerAllReduce:
    """All-reduce of small fp32 buffers over peer-mapped symmetric memory (rq_peer_allreduce_f32): one kernel
    launch on the caller's stream, NVLink loads, no NCCL on the data path, capturable into the step's CUDA graph.

    `buffer` is a [slots, *shape] fp32 tensor in a symmetric allocation: the producer (e.g. the fused backward,
    through `functional.quantize_backward(..., g_codebooks_out=peer.buffer[s])`) writes slot s, then
    `all_reduce_(s)` replaces it with the sum (or mean) over the ranks, bit-identical on every rank.  All ranks
    must make the same calls in the same order.  torch.distributed._symmetric_memory supplies allocation,
    rendezvous and the peer pointer tables."""

    def __init__(self, shape, slots: int = 1, group=None, device=None):
        import math
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from . import _lib as L
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("PeerAllReduce needs an initialised process group")
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.shape = tuple(int(v) for v in shape)
        self.numel = math.prod(self.shape)
        lib = L.lib()
        if self.numel % 4 or self.numel > lib.rq_peer_allreduce_max_floats():
            raise ValueError(f"PeerAllReduce: {self.numel} floats per slot (need a multiple of 4, at most "
                             f"{lib.rq_peer_allreduce_max_floats()})")
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.buffer = symm.empty((int(slots),) + self.shape, dtype=torch.float32, device=device)
        self.handle = symm.rendezvous(self.buffer, self.group)
        if self.handle.signal_pad_size < lib.rq_peer_allreduce_pad_bytes():
            raise RuntimeError("PeerAllReduce: signal pad too small")
        self.buffer.zero_()
        self._epoch = torch.zeros(lib.rq_peer_allreduce_state_words(), dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        dist.barrier(self.group)                         # every rank's buffer, pad and counters are in place

    @staticmethod
    def available(group=None) -> bool:
        import torch.distributed as dist
        try:
            import torch.distributed._symmetric_memory  # noqa: F401
        except Exception:
            return False
        return (dist.is_available() and dist.is_initialized() and dist.get_backend(group) == "nccl"
                and torch.cuda.is_available() and dist.get_world_size(group) <= 16)

    def all_reduce_(self, slot: int = 0, average: bool = True, out=None, stream=None):
        """buffer[slot] <- sum (mean) over ranks of buffer[slot]; `out` redirects the result to another fp32
        tensor of the same size (the slot then keeps this rank's own contribution).  `stream` (a
        torch.cuda.Stream; default: the current one) is where the kernel is launched."""
        from . import _lib as L
        dst = self.buffer[slot] if out is None else out
        if dst.numel() != self.numel or dst.dtype != torch.float32 or not dst.is_contiguous() or not dst.is_cuda:
            raise ValueError("PeerAllReduce.all_reduce_: out must be a contiguous CUDA fp32 tensor of the slot's size")
        rc = L.lib().rq_peer_allreduce_f32(self.handle.buffer_ptrs_dev, self.handle.signal_pad_ptrs_dev, self.rank,
                                           self.world, int(slot) * self.numel, self.numel, dst.data_ptr(),
                                           (1.0 / self.world) if average else 1.0, self._epoch.data_ptr(),
                                           L.stream_of(self.buffer) if stream is None else stream.cuda_stream)
        L.check(rc, "rq_peer_allreduce_f32")
        return dst

    def check(self) -> None:
        """Raises if a barrier of an earlier `all_reduce_` on this device timed out (a rank more than a few seconds
        behind: rank-0-only evaluation or checkpointing in the middle of a step).  The affected output was filled
        with NaN by the kernel, never with a partial sum.  Synchronises the device: call it at a step boundary (e.g.
        next to optimizer.step())."""
        from . import _lib as L
        if L.lib().rq_debug_peer_timeout_flag():
            raise RuntimeError("PeerAllReduce: a rank did not reach the exchange within the timeout; the reduced "
                               "buffer was poisoned with NaN (ranks must enter all_reduce_ within a few seconds of "
                               "each other)")


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


def sharded_corpus_table(local_ids: torch.Tensor, n: int, codebook_size: int, group=None,
                         max_dense_keys: int = 1 << 26, table_fn=None) -> torch.Tensor:
    """[N/G, L] ids of this rank's contiguous row shard (shard_range order) -> the full [N, L+1] corpus table
    `[ids | dedup]` on every rank (semids.py:86-99: dedup = number of EARLIER corpus items with the same tuple).

    SURVEY.md 8(e) partitioning: dedup(item) = rank of the item among this shard's items with the same key (local
    sort, rq_corpus_table) + number of items with that key on the LOWER ranks.  The second term comes from one
    all-gather of the per-rank key-count table (radix^L int32 entries: 64 MB at 3 x 256) — no rank sorts more than
    its own N/G keys.  Key spaces beyond `max_dense_keys` entries fall back to gathering the ids and sorting the
    whole corpus on every rank."""
    import torch.distributed as dist
    if table_fn is None:                                 # the CUDA dedup kernel; tests inject a CPU stand-in
        from . import functional as RF
        table_fn = RF.corpus_table
    if not is_distributed(group):
        return table_fn(local_ids, codebook_size)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    Lv = local_ids.shape[1]
    n_keys = int(codebook_size) ** Lv
    if n_keys > max_dense_keys:
        ids = gather_row_shards(local_ids, n, group)
        return table_fn(ids, codebook_size)
    local = table_fn(local_ids, codebook_size) if local_ids.shape[0] else \
        torch.zeros((0, Lv + 1), dtype=torch.int64, device=local_ids.device)
    key = torch.zeros(local_ids.shape[0], dtype=torch.int64, device=local_ids.device)
    for l in range(Lv):
        key = key * int(codebook_size) + local_ids[:, l]
    counts = torch.bincount(key, minlength=n_keys).to(torch.int32)
    flat = torch.empty((world * n_keys,), dtype=torch.int32, device=local_ids.device)
    dist.all_gather_into_tensor(flat, counts, group=group)
    all_counts = flat.view(world, n_keys)
    if rank > 0:
        below = all_counts[:rank].sum(dim=0, dtype=torch.int64)
        local[:, -1] += below[key]
    del all_counts, flat
    return gather_row_shards(local.contiguous(), n, group)
