"""Lloyd k-means for codebook initialisation behind the reference's `Kmeans` interface.

Mirrors init/kmeans.py of the reference: `kmeans_init_(tensor, x)`, `Kmeans(k, max_iters,
stop_threshold).run(x) -> KmeansOutput(centroids, assignment)`.  The assignment + accumulation
(:41-57) is one CUDA launch per iteration; the host reads back three scalars per iteration
(max shift, number of empty clusters, number of changed assignments) instead of the K + 1 host
syncs of the reference's Python loop.

Rank-local by default, like the reference: it runs k-means independently per rank on the same first
20,000 rows (train_rqvae.py:140) and never synchronises, so `Kmeans(k).run(x)` makes no collective call even when
torch.distributed is initialised.

Multi-GPU is OPT-IN — `Kmeans(k, group=some_group)` or `distributed=True` — and then every rank of the group
must call `run` with ITS row shard of the global point set: points are row-sharded, centroids replicated; per
iteration the [K, D] sums, the [K] counts and the changed-assignment count are all-reduced (NCCL over NVLink);
initial rows and empty-cluster re-seeds are chosen by global row index on rank 0 and broadcast, so every rank
ends up with bit-identical centroids.
"""
from typing import NamedTuple, Optional

import numpy as np
import torch

from .. import functional as RF


def kmeans_init_(tensor: torch.Tensor, x: torch.Tensor, group=None):
    assert tensor.dim() == 2
    assert x.dim() == 2
    with torch.no_grad():
        k, _ = tensor.shape
        out = Kmeans(k=k, group=group).run(x)
        tensor.data.copy_(out.centroids)


class KmeansOutput(NamedTuple):
    centroids: torch.Tensor
    assignment: torch.Tensor


def _dist_ready(group) -> bool:
    import torch.distributed as dist
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1


class Kmeans:
    def __init__(self, k: int, max_iters: Optional[int] = None, stop_threshold: float = 1e-10,
                 group=None, distributed: Optional[bool] = None) -> None:
        self.k = k
        self.iters = max_iters
        self.stop_threshold = stop_threshold
        self.centroids = None
        self.assignment = None
        self.group = group
        # opt-in: sharded iff a group is passed or distributed=True (and the group has more than one rank)
        self.distributed = distributed
        self.n_iters_run = 0

    # -- sharding helpers ----------------------------------------------------------------
    def _is_dist(self) -> bool:
        want = (self.group is not None) if self.distributed is None else bool(self.distributed)
        return want and _dist_ready(self.group)

    def _shard_layout(self, n_local: int, device):
        """(offset of this rank's rows in the global order, global row count)."""
        if not self._is_dist():
            return 0, n_local
        import torch.distributed as dist
        ws, rk = dist.get_world_size(self.group), dist.get_rank(self.group)
        sizes = torch.zeros(ws, dtype=torch.int64, device=device)
        sizes[rk] = n_local
        dist.all_reduce(sizes, group=self.group)
        sizes = sizes.cpu()
        return int(sizes[:rk].sum()), int(sizes.sum())

    def _global_rows(self, x: torch.Tensor, rows, offset: int) -> torch.Tensor:
        """Rows of the GLOBAL point set by global index; the owning rank contributes, others add 0."""
        rows = torch.as_tensor(np.asarray(rows), dtype=torch.int64)
        local = rows - offset
        mine = (local >= 0) & (local < x.shape[0])
        out = torch.zeros((rows.numel(), x.shape[1]), device=x.device, dtype=x.dtype)
        if bool(mine.any()):
            out[mine.to(x.device)] = x[local[mine].to(x.device)]
        if self._is_dist():
            import torch.distributed as dist
            dist.all_reduce(out, group=self.group)
        return out

    # -- reference interface ---------------------------------------------------------------
    def _init_centroids(self, x: torch.Tensor) -> None:
        """K distinct random rows from the global NumPy RNG (kmeans.py:34-38)."""
        self._offset, self._n_global = self._shard_layout(x.shape[0], x.device)
        init_idx = np.random.choice(self._n_global, self.k, replace=False)
        if self._is_dist():
            # one draw for everybody: rank 0's indices win
            import torch.distributed as dist
            t = torch.as_tensor(init_idx, dtype=torch.int64, device=x.device)
            dist.broadcast(t, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0,
                           group=self.group)
            init_idx = t.cpu().numpy()
        self.centroids = self._global_rows(x, init_idx, self._offset).contiguous()
        self.assignment = None

    def _update_centroids(self, x: torch.Tensor):
        """One iteration (kmeans.py:40-58).  Returns (max_shift, n_changed, n_empty) as Python numbers."""
        if self._n_global == 0:
            raise ValueError("Can not choose random element from x, x is empty")      # kmeans.py:55
        assign, sums, counts, n_changed = RF.kmeans_assign_accumulate(x, self.centroids, self.assignment)
        if self._is_dist():
            import torch.distributed as dist
            packed = torch.cat([sums.reshape(-1).double(), counts.double(), n_changed.double()])
            dist.all_reduce(packed, group=self.group)
            kd = sums.numel()
            sums = packed[:kd].float().reshape(sums.shape)
            counts = packed[kd:kd + counts.numel()].round().long()
            n_changed = packed[-1:].round().long()
        max_shift, n_empty, empty_ids = RF.kmeans_update(sums, counts, self.centroids)
        self.assignment = assign
        stats = torch.cat([max_shift.double(), n_empty.double(), n_changed.double()]).cpu()   # the one host sync
        shift, ne, nc = float(stats[0]), int(stats[1]), int(stats[2])
        if ne > 0:
            # empty cluster -> a random row of x, one torch.randint per cluster in cluster order
            # (kmeans.py:51-53).  Rows are global indices so all ranks agree.
            rows = [int(torch.randint(0, self._n_global, (1,))) for _ in range(ne)]
            if self._is_dist():
                import torch.distributed as dist
                t = torch.as_tensor(rows, dtype=torch.int64, device=x.device)
                dist.broadcast(t, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0,
                               group=self.group)
                rows = t.cpu().tolist()
            new_rows = self._global_rows(x, rows, self._offset)
            ids = empty_ids[:ne]
            old = self.centroids[ids]
            self.centroids[ids] = new_rows
            shift = max(shift, float(torch.norm(new_rows - old, dim=1).max()))
        return shift, nc, ne

    def run(self, x: torch.Tensor) -> KmeansOutput:
        x = x.detach()
        self._init_centroids(x)
        i = 0
        while self.iters is None or i < self.iters:
            shift, n_changed, n_empty = self._update_centroids(x)
            self.n_iters_run = i + 1
            # kmeans.py:67 stops when no centroid moved (1e-10 in fp32 = bit-identical sums).  Float
            # reductions are not order-stable, so "no assignment changed" — which makes the
            # reference's shift exactly 0 — is accepted as the same condition.
            if shift < self.stop_threshold or (n_changed == 0 and n_empty == 0):
                break
            i += 1
        return KmeansOutput(centroids=self.centroids, assignment=self.assignment)
