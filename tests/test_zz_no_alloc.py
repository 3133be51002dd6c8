"""GPU: rq_id_diversity_static, the entry point that allocates nothing (bitmap in a file-scope device array), against
rq_id_diversity with a caller workspace, through the C-ABI.  Collected last on purpose
(file name)."""
import pytest
import torch

from rq_vae_recommender_b200 import _lib as L
from rq_vae_recommender_b200 import functional as RF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ENTROPY_RTOL = 2e-5          # the tolerance test_id_diversity_vs_oracle uses (double atomics, block order varies)


def _table(N, n_levels, K, seed, narrow):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(0, K, (N, n_levels), generator=g)
    if narrow:                                     # few distinct ids: stale bits of an earlier call would show
        ids = ids % 3
    ids[N // 2:N // 2 + 40] = ids[0]
    return RF.corpus_table(ids.to(DEV), K)


def _static_call(table, K, entropy, used, max_dup):
    N, W = table.shape
    with torch.cuda.device(table.device):
        rc = L.lib().rq_id_diversity_static(L.ptr(table), N, W - 1, int(K), L.ptr(entropy), L.ptr(used), L.ptr(max_dup),
                                            L.stream_of(table))
    return rc


@pytest.mark.parametrize("n_levels,K", [(3, 256), (3, 100000)])      # shared-memory bitmap / global bitmap (37.5 KB)
def test_static_matches_workspace_entry_two_calls(n_levels, K):
    assert int(L.lib().rq_id_diversity_static_max_bytes()) == 65536
    tables = [_table(60000, n_levels, K, 11, narrow=False), _table(20000, n_levels, K, 12, narrow=True)]
    want = [RF.id_diversity(t, K) for t in tables]
    entropy = torch.empty((), device=DEV, dtype=torch.float64)
    used = torch.empty((n_levels,), device=DEV, dtype=torch.int64)
    max_dup = torch.empty((), device=DEV, dtype=torch.int64)
    torch.cuda.synchronize()
    for t, (w_entropy, w_used, w_max_dup) in zip(tables, want):    # two calls in a row on the shared array
        before = torch.cuda.memory_allocated()
        rc = _static_call(t, K, entropy, used, max_dup)
        L.check(rc, "rq_id_diversity_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert used.tolist() == w_used.tolist()
        assert int(max_dup) == int(w_max_dup)
        assert abs(float(entropy) - float(w_entropy)) <= ENTROPY_RTOL * max(1.0, abs(float(w_entropy)))


def test_static_rejects_sizes_beyond_the_array():
    table = torch.zeros((1000, 5), device=DEV, dtype=torch.int64)    # 4 * 6250 words * 4 B = 100,000 B > 65,536
    entropy = torch.zeros((), device=DEV, dtype=torch.float64)
    used = torch.zeros((4,), device=DEV, dtype=torch.int64)
    max_dup = torch.zeros((), device=DEV, dtype=torch.int64)
    rc = _static_call(table, 200000, entropy, used, max_dup)
    assert rc == -4                                                  # RQ_ERR_WORKSPACE_TOO_SMALL
    msg = L.lib().rq_last_error().decode()
    assert "exceeds" in msg and "65536" in msg
    with pytest.raises(L.RqEngineError):
        L.check(rc, "rq_id_diversity_static")
