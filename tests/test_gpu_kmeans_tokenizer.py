"""GPU parity: k-means and the tokenizer integer paths vs golden fixtures and the CPU oracle."""
import numpy as np
import pytest
import torch

import rq_oracle as O
import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import functional as RF
from conftest import load_golden
from parity_utils import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_kmeans_single_step_golden():
    z = load_golden("kmeans_small.npz")
    x = torch.from_numpy(z["x"]).to(DEV)
    c = torch.from_numpy(z["step_c0"]).to(DEV).clone()
    assign, sums, counts, n_changed = RF.kmeans_assign_accumulate(x, c)
    assert np.array_equal(assign.cpu().numpy(), z["step_assign"])
    assert int(n_changed) == x.shape[0] and int(counts.sum()) == x.shape[0]
    max_shift, n_empty, _ = RF.kmeans_update(sums, counts, c)
    assert int(n_empty) == 0
    assert_close(c.cpu(), torch.from_numpy(z["step_c1"]), "centroids", tol=1e-5)
    want_shift = np.linalg.norm(z["step_c1"] - z["step_c0"], axis=1).max()
    assert abs(float(max_shift) - want_shift) <= 1e-4 * want_shift
    # second pass from the same assignment: only genuinely moved rows are counted
    assign2, _, _, n_changed2 = RF.kmeans_assign_accumulate(x, torch.from_numpy(z["step_c0"]).to(DEV), assign)
    assert int(n_changed2) == 0


def test_kmeans_empty_clusters_golden():
    z = load_golden("kmeans_small.npz")
    x = torch.from_numpy(z["x"]).to(DEV)
    c = torch.from_numpy(z["empty_c0"]).to(DEV).clone()
    assign, sums, counts, _ = RF.kmeans_assign_accumulate(x, c)
    assert np.array_equal(assign.cpu().numpy(), z["empty_assign"])
    _, n_empty, empty_ids = RF.kmeans_update(sums, counts, c)
    assert int(n_empty) == 2 and empty_ids[:2].tolist() == [2, 9]
    # host re-seed with the rows the reference drew (kmeans.py:53)
    c[empty_ids[:2]] = x[torch.from_numpy(z["empty_reseed_rows"]).to(DEV)]
    assert_close(c.cpu(), torch.from_numpy(z["empty_c1"]), "centroids after reseed", tol=1e-5)


def test_kmeans_full_run_golden_and_init():
    z = load_golden("kmeans_small.npz")
    x = torch.from_numpy(z["x"]).to(DEV)
    np.random.seed(8)                            # the fixture's seed: same np.random.choice draw (kmeans.py:36)
    out = R.Kmeans(k=16).run(x)
    assert out.centroids.shape == (16, 8) and out.assignment.dtype == torch.int64
    assert np.array_equal(out.assignment.cpu().numpy(), z["run_assignment"])
    assert_close(out.centroids.cpu(), torch.from_numpy(z["run_centroids"]), "run centroids", tol=1e-5)
    w = torch.zeros(16, 8, device=DEV)
    np.random.seed(8)
    R.kmeans_init_(w, x)
    assert_close(w.cpu(), torch.from_numpy(z["init_tensor"]), "kmeans_init_", tol=1e-5)
    with pytest.raises(ValueError):
        R.Kmeans(k=4).run(torch.zeros(0, 8, device=DEV))     # empty x (np.random.choice / kmeans.py:55)


@pytest.mark.parametrize("B,D,K", [(20000, 32, 256), (5000, 64, 100), (3001, 16, 1024)])
def test_kmeans_step_vs_oracle(B, D, K):
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, D, generator=g) + 2.0 * torch.randn(8, D, generator=g)[torch.randint(0, 8, (B,), generator=g)]
    c0 = x[torch.randperm(B, generator=g)[:K]].clone()
    c1, a1 = O.kmeans_update(x, c0, reseed_rows=[0] * K)
    cd = c0.to(DEV).clone()
    assign, sums, counts, _ = RF.kmeans_assign_accumulate(x.to(DEV), cd)
    mism = (assign.cpu() != a1)
    # direct-form distances: ties only at fp32 rounding level
    assert mism.float().mean() < 1e-4
    if not bool(mism.any()):
        _, n_empty, empty_ids = RF.kmeans_update(sums, counts, cd)
        keep = torch.ones(K, dtype=torch.bool)
        keep[empty_ids[:int(n_empty)].cpu()] = False
        assert_close(cd.cpu()[keep], c1[keep], "centroids", tol=1e-5)
    # invariants at any size: counts and sums are consistent with the assignment
    assert int(counts.sum()) == B
    want = torch.zeros(K, D, dtype=torch.float64).index_add_(0, assign.cpu(), x.double())
    assert_close(sums.cpu(), want, "sums", tol=1e-5)


@pytest.mark.parametrize("B,K,D", [(4096, 256, 32), (100003, 256, 32), (50000, 77, 32), (60001, 256, 64), (9000, 100, 64)])
def test_kmeans_tensor_core_assignment(B, K, D):
    """impl=TCGEN05 (the fused quantiser's distance kernel with L = 1) against the exact SIMT k-means kernel:
    assignments may differ only at fp32-rounding ties; sums / counts / n_changed follow the assignment."""
    g = torch.Generator().manual_seed(B + K)
    x = (torch.randn(B, D, generator=g) + 2.0 * torch.randn(8, D, generator=g)[torch.randint(0, 8, (B,), generator=g)]).to(DEV)
    c = x[torch.randperm(B, generator=g)[:K].to(DEV)].clone()
    a_s, sums_s, counts_s, _ = RF.kmeans_assign_accumulate(x, c, impl=RF.IMPL_SIMT)
    a_t, sums_t, counts_t, nch = RF.kmeans_assign_accumulate(x, c, impl=RF.IMPL_TCGEN05)
    assert int(nch) == B
    mism = a_s != a_t
    assert float(mism.float().mean()) < 1e-4
    if bool(mism.any()):                       # a differing row must be a genuine near-tie
        xm = x[mism].double()
        d = ((xm[:, None, :] - c.double()[None]) ** 2).sum(-1)
        ds, dt = d.gather(1, a_s[mism][:, None]), d.gather(1, a_t[mism][:, None])
        assert bool(((dt - ds).abs() <= 1e-5 * (xm ** 2).sum(1, keepdim=True).clamp_min(1.0)).all())
    assert int(counts_t.sum()) == B
    want = torch.zeros(K, D, dtype=torch.float64, device=DEV).index_add_(0, a_t, x.double())
    assert_close(sums_t.cpu(), want.cpu(), "sums", tol=1e-5)
    assert torch.equal(counts_t, torch.bincount(a_t, minlength=K))
    # second pass from the stored assignment: nothing changes; perturbing a few entries is counted exactly
    a2 = a_t.clone()
    _, _, _, n0 = RF.kmeans_assign_accumulate(x, c, a2, impl=RF.IMPL_TCGEN05)
    assert int(n0) == 0
    a3 = a_t.clone()
    a3[:5] = (a3[:5] + 1) % K
    _, _, _, n5 = RF.kmeans_assign_accumulate(x, c, a3, impl=RF.IMPL_TCGEN05)
    assert int(n5) == 5 and torch.equal(a3, a_t)


def test_tokenizer_golden_end_to_end():
    z = load_golden("tokenizer_small.npz")
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")}
    tok = R.SemanticIdTokenizer(input_dim=24, output_dim=8, hidden_dims=[16], codebook_size=6, n_layers=3,
                                n_cat_feats=0)
    tok.rq_vae.load_state_dict(sd)
    tok = tok.to(DEV)
    tok.corpus_batch_size = 512
    x = torch.from_numpy(z["x"])

    class Items(torch.utils.data.Dataset):
        def __len__(self):
            return x.shape[0]

        def __getitem__(self, idx):
            ids = torch.tensor(idx).unsqueeze(0)
            return R.SeqBatch(user_ids=-torch.ones_like(ids.squeeze(0)), ids=ids, ids_fut=-torch.ones_like(ids.squeeze(0)),
                              x=x[idx], x_fut=-torch.ones_like(ids.squeeze(0)), seq_mask=torch.ones_like(ids, dtype=bool))

    table = tok.precompute_corpus_ids(Items())
    assert table.dtype == torch.int64 and table.shape == (1300, 4)
    assert np.array_equal(table.cpu().numpy(), z["cached_ids"])          # bit-exact incl. dedup column
    for p in (1, 2, 3, 4):
        q = torch.from_numpy(z[f"prefix_q{p}"]).to(DEV)
        tok.strict_reference_quirk = True
        assert np.array_equal(tok.exists_prefix(q).cpu().numpy(), z[f"prefix_r{p}"])
        tok.strict_reference_quirk = False
        full = tok.exists_prefix(q).cpu().numpy()
        assert np.array_equal(full, O.exists_prefix(z["cached_ids"], z[f"prefix_q{p}"]))
        assert full[36, 4]
    batch = R.SeqBatch(user_ids=torch.arange(6, device=DEV), ids=torch.from_numpy(z["fw_ids"]).to(DEV),
                       ids_fut=torch.from_numpy(z["fw_ids_fut"]).to(DEV), x=torch.zeros(6, 5, 24, device=DEV),
                       x_fut=torch.zeros(6, 1, 24, device=DEV), seq_mask=torch.from_numpy(z["fw_seq_mask"]).to(DEV))
    tk = tok(batch)
    assert np.array_equal(tk.sem_ids.cpu().numpy(), z["fw_sem_ids"])
    assert np.array_equal(tk.sem_ids_fut.cpu().numpy(), z["fw_sem_ids_fut"])
    assert np.array_equal(tk.seq_mask.cpu().numpy(), z["fw_seq_mask_out"])
    assert np.array_equal(tk.token_type_ids.cpu().numpy(), z["fw_token_type_ids"])
    assert np.array_equal(tk.token_type_ids_fut.cpu().numpy(), z["fw_token_type_ids_fut"])


@pytest.mark.parametrize("N,L,K", [(0, 3, 256), (1, 3, 256), (100000, 3, 8), (250000, 4, 1024), (5000, 2, 3)])
def test_corpus_table_vs_oracle(N, L, K):
    g = torch.Generator().manual_seed(N + L)
    ids = torch.randint(0, K, (N, L), generator=g)
    if N > 10:
        ids[N // 2:N // 2 + 5] = ids[0]           # forced duplicates far apart
    table = RF.corpus_table(ids.to(DEV), K)
    assert np.array_equal(table.cpu().numpy(), O.corpus_ids_table(ids.numpy()).reshape(N, L + 1))
    # the [B,L] view over [L,B] memory the fused kernel hands out is accepted as-is
    view = ids.t().contiguous().to(DEV).t()
    assert np.array_equal(RF.corpus_table(view, K).cpu().numpy(), table.cpu().numpy())


@pytest.mark.parametrize("N,L,K", [(1, 3, 256), (100000, 3, 8), (250000, 4, 1024), (5000, 2, 3), (40000, 3, 100000)])
def test_id_diversity_vs_oracle(N, L, K):
    """rq_id_diversity against the torch.unique form of train_rqvae.py:227-237 (oracle.id_diversity): usage counts
    and max duplicates exact, entropy to float32 rounding of the reference's own sum."""
    g = torch.Generator().manual_seed(7 * N + L)
    ids = torch.randint(0, K, (N, L), generator=g)
    if N > 10:
        ids[N // 2:N // 2 + 50] = ids[0]          # one heavy tuple
        ids[:, L - 1] = ids[:, L - 1] % max(K // 2, 1)   # an under-used last level
    table = RF.corpus_table(ids.to(DEV), K)
    want = O.id_diversity(table.cpu(), L, K)
    entropy, used, max_dup = RF.id_diversity(table, K)
    for cid in range(L):
        assert int(used[cid]) == round(want[f"codebook_usage_{cid}"] * K)
    assert int(max_dup) == int(table[:, -1].max())
    assert abs(float(entropy) - want["rqvae_entropy"]) <= 2e-5 * max(1.0, abs(want["rqvae_entropy"]))


def test_tokenizer_id_diversity_log_keys():
    """SemanticIdTokenizer.id_diversity returns the reference's log keys and values (train_rqvae.py:234-239)."""
    tok = R.SemanticIdTokenizer(input_dim=24, output_dim=8, hidden_dims=[16], codebook_size=6, n_layers=3, n_cat_feats=0)
    with pytest.raises(Exception):
        tok.id_diversity()
    ids = torch.randint(0, 6, (1300, 3), generator=torch.Generator().manual_seed(5))
    tok.cached_ids = RF.corpus_table(ids.to(DEV), 6)
    got = tok.id_diversity()
    want = O.id_diversity(tok.cached_ids.cpu(), 3, 6)
    assert set(got) == set(want)
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-5 * max(1.0, abs(want[k])), k


def test_c_abi_host_entry_point():
    """rq_semantic_ids_host: host buffers in, host buffers out, copies inside."""
    import ctypes
    from rq_vae_recommender_b200 import _lib
    B, D, K, L = 3000, 32, 256, 3
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, D, generator=g)
    cbs = torch.randn(L, K, D, generator=g)
    cb_dev = cbs.to(DEV)
    ids = np.empty((L, B), dtype=np.int64)
    loss = np.empty((B,), dtype=np.float32)
    xn = x.numpy()
    torch.cuda.synchronize()
    rc = _lib.lib().rq_semantic_ids_host(xn.ctypes.data, cb_dev.data_ptr(), B, D, K, L, 1, ids.ctypes.data,
                                         loss.ctypes.data, 0, None)
    _lib.check(rc)
    ref = O.rq_forward(x, list(cbs), O.MODE_EVAL)
    assert (torch.from_numpy(ids) == ref.sem_ids).float().mean() > 0.999
    ok = (torch.from_numpy(ids) == ref.sem_ids).all(0)
    assert_close(torch.from_numpy(loss)[ok], ref.quantize_loss[ok], "loss")


def test_c_abi_host_entry_point_chunked_pipeline(monkeypatch):
    """The host entry pipelines row chunks over three streams; a ragged chunking (7 chunks, short
    tail, pinned and pageable sources) must give exactly the ids and loss of the one-shot device call."""
    from rq_vae_recommender_b200 import _lib
    B, D, K, L = 6500, 32, 256, 3
    g = torch.Generator().manual_seed(15)
    x = torch.randn(B, D, generator=g)
    cb_dev = torch.randn(L, K, D, generator=g).to(DEV)
    _, _, ids_ref, loss_ref = RF.quantize_forward(x.to(DEV), cb_dev, RF.MODE_EVAL)
    monkeypatch.setenv("RQ_HOST_CHUNK_ROWS", "1000")
    for src in (x, x.pin_memory()):
        ids = torch.full((L, B), -1, dtype=torch.int64).pin_memory()
        loss = torch.full((B,), -1.0).pin_memory()
        rc = _lib.lib().rq_semantic_ids_host(src.data_ptr(), cb_dev.data_ptr(), B, D, K, L, 1, ids.data_ptr(),
                                             loss.data_ptr(), 0, None)
        _lib.check(rc)
        assert torch.equal(ids, ids_ref.cpu())
        assert torch.equal(loss, loss_ref.cpu())


def test_device_prefetcher_orders_and_completes():
    """DevicePrefetcher (data/utils.py) returns the batches in order, complete when the consumer
    stream reads them, and stops with the source."""
    from rq_vae_recommender_b200.data import utils as U
    host = [torch.full((1 << 20,), float(i)).pin_memory() for i in range(5)]
    got = []
    for t in U.DevicePrefetcher(host, DEV, depth=2):
        assert t.is_cuda
        got.append(float(t.sum().item()) / (1 << 20))
    assert got == [0.0, 1.0, 2.0, 3.0, 4.0]
    it = U.DevicePrefetcher(U.cycle(host[:2]), DEV)
    assert [int(next(it)[0].item()) for _ in range(5)] == [0, 1, 0, 1, 0]
    # static ring (`into`): the batches land in the given buffers in turn, and a buffer is not overwritten
    # before the consumer stream's earlier reads of it (a long-running reduction here) have run
    bufs = [torch.empty(1 << 20, device=DEV) for _ in range(2)]
    it = U.DevicePrefetcher(U.cycle(host), DEV, into=bufs)
    sums = []
    for i in range(7):
        t = next(it)
        assert t is bufs[i % 2]
        acc = t.double()
        for _ in range(20):
            acc = acc * 1.0 + 0.0
        sums.append(acc.sum())
    assert [float(v.item()) / (1 << 20) for v in sums] == [float(i % 5) for i in range(7)]
    with pytest.raises(ValueError):
        U.DevicePrefetcher(host, DEV, depth=2, into=bufs)


def test_full_size_properties_c3_kmeans_10m():
    """BASELINE config 3 size on one GPU (10M x 32 latents, 256 centroids): size-independent invariants of
    one assignment + accumulation pass and its update (no oracle run at this size)."""
    n, D, K = 10_000_000, 32, 256
    g = torch.Generator(device=DEV).manual_seed(3)
    x = torch.randn(n, D, device=DEV, generator=g)
    c = x[:K].clone()
    assign, sums, counts, nch = RF.kmeans_assign_accumulate(x, c)
    assert int(nch) == n and int(counts.sum()) == n
    assert int(assign.min()) >= 0 and int(assign.max()) < K
    assert torch.equal(counts, torch.bincount(assign, minlength=K))
    want = torch.zeros(K, D, dtype=torch.float64, device=DEV).index_add_(0, assign, x.double())
    assert_close(sums, want.float(), "sums", tol=2e-5)
    # the chosen centroid is the nearest one (direct form, float64) on a 64k-row sample, up to fp32 ties
    idx = torch.arange(0, n, n // 65536, device=DEV)[:65536]
    d = torch.cdist(x[idx].double(), c.double()) ** 2
    excess = d.gather(1, assign[idx][:, None]).squeeze(1) - d.min(1).values
    assert bool((excess <= 1e-5 * ((x[idx].double() ** 2).sum(1) + 1.0)).all())
    # idempotence: a second pass from the stored assignment changes nothing
    _, sums2, counts2, n0 = RF.kmeans_assign_accumulate(x, c, assign)
    assert int(n0) == 0 and torch.equal(counts2, counts)
    assert_close(sums2, sums, "sums (second pass)", tol=2e-5)
    # update: centroid = mean of its points; total mass is conserved
    c2 = c.clone()
    RF.kmeans_update(sums, counts, c2)
    nz = counts > 0
    assert_close(c2[nz], (want[nz] / counts[nz][:, None]).float(), "centroids", tol=2e-5)


def test_full_size_properties_c4_tokenize_d64():
    """BASELINE config 4 shape (ml32m: d=64, 3x256), one 4M-item chunk of the 100M corpus: ids-only
    tokenisation is deterministic, in range, equal to the exact SIMT family on a slice, and every id is
    the nearest code of the residual it was chosen for (float64) up to the tie band."""
    n, D, K, L = 4 << 20, 64, 256, 3
    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randn(n, D, device=DEV, generator=g)
    cb = torch.stack([x[l * K:(l + 1) * K] * (0.6 ** l) for l in range(L)]).contiguous()
    _, _, ids, _ = RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False)
    _, _, ids2, _ = RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False)
    assert torch.equal(ids, ids2)
    assert int(ids.min()) >= 0 and int(ids.max()) < K
    sl = slice(1_000_000, 1_000_000 + 50_000)
    _, res_s, ids_s, _ = RF.quantize_forward(x[sl].contiguous(), cb, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    assert torch.equal(ids[:, sl], ids_s)
    for l in range(L):
        r = res_s[l][:8192].double()
        d = torch.cdist(r, cb[l].double()) ** 2
        excess = d.gather(1, ids_s[l][:8192, None]).squeeze(1) - d.min(1).values
        scale = (r ** 2).sum(1) + (cb[l].double() ** 2).sum(1).max()
        assert bool((excess <= 1e-6 * scale).all())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (run through gpurun --gpus 2)")
def test_peer_allreduce_two_ranks():
    """parallel.PeerAllReduce / rq_peer_allreduce_f32 on two ranks: sums equal NCCL's, identical on both ranks, in place
    and out of place, eager and replayed from a CUDA graph (tools/test_peer_allreduce.py under torchrun)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(root, "tools", "test_peer_allreduce.py")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ok=True timeout_flag=0" in r.stdout



def test_kmeans_fixed_point_accumulation_outliers_and_specials():
    """rq_kmeans_accumulate_fx_kernel: components far outside the fixed-point range (outliers 1e7 x the centroid scale),
    +-Inf and NaN take the floating-point path, so sums equal a float64 index_add where that is finite and propagate
    Inf / NaN exactly where it is not; tiny components do not vanish; a row count that is not a multiple of 8."""
    g = torch.Generator().manual_seed(3)
    B, D, K = 50003, 32, 256
    x = torch.randn(B, D, generator=g)
    x[::97] *= 1e7                                  # outliers
    x[5::1013] *= 1e-6                              # far below one fixed-point step: must still be accumulated to within it
    c = torch.randn(K, D, generator=g)
    xd, cd = x.to(DEV), c.to(DEV)
    a, sums, counts, _ = RF.kmeans_assign_accumulate(xd, cd, impl=RF.IMPL_TCGEN05)
    want = torch.zeros(K, D, dtype=torch.float64, device=DEV).index_add_(0, a, xd.double())
    assert torch.equal(counts, torch.bincount(a, minlength=K))
    assert_close(sums.cpu(), want.cpu(), "sums with outliers", tol=1e-5)
    x2 = x.clone()
    x2[11, 3] = float("inf"); x2[12, 4] = float("-inf"); x2[13, 5] = float("nan")
    xd2 = x2.to(DEV)
    a2, sums2, counts2, _ = RF.kmeans_assign_accumulate(xd2, cd, a.clone(), impl=RF.IMPL_TCGEN05)
    want2 = torch.zeros(K, D, dtype=torch.float64, device=DEV).index_add_(0, a2, xd2.double())
    fin = torch.isfinite(want2)
    assert torch.equal(torch.isnan(sums2), torch.isnan(want2))
    assert torch.equal(torch.isinf(sums2), torch.isinf(want2))
    assert_close(torch.where(fin, sums2.double(), torch.zeros_like(want2)).cpu(),
                 torch.where(fin, want2, torch.zeros_like(want2)).cpu(), "finite sums", tol=1e-5)
