"""Side numbers for the other BASELINE configs (single GPU, CUDA events, inputs resident in HBM):
 C4-like  ids-only tokenisation throughput on a 4M-item latent chunk (d=32 tcgen05 family; d=64 SIMT family)
 C3-like  one k-means level-iteration (assign + accumulate + update) over 10M x 32 latents, K = 256
 C5-like  Gumbel-softmax / rotation-trick forward+backward at 4 x 1024, d = 64 (SIMT family)
Prints one JSON object; `python tools/bench_side.py > profiles/r1_side_numbers.json` on a B200."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")
HBM = 6551.0
if os.path.exists("MEASURED_PEAKS.json"):
    HBM = float(json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"])


def timed(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def codebooks(L, K, D, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.stack([torch.randn(K, D, generator=g) * (0.6 ** l) for l in range(L)]).to(dev)


out = {}
# ---- C4-like: ids-only over a resident chunk -----------------------------------------------------------
for D, n in ((32, 4 << 20), (64, 4 << 20)):
    x = torch.randn(n, D, device=dev)
    cb = codebooks(3, 256, D)
    ms = timed(lambda: RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                           want_loss=False), 5)
    byts = n * (4 * D + 8 * 3)
    out[f"tokenize_ids_only_d{D}_3x256"] = {
        "items": n, "ms": ms, "items_per_s": n / (ms * 1e-3), "algorithmic_GBps": byts / (ms * 1e-3) / 1e9,
        "tcgen05": bool(R.lib().rq_tcgen05_supported(D, 256, 3, RF.MODE_EVAL, 1))}
    del x
# ---- C4 protocol: a resident 4M-item d=64 chunk cycled 25x, ids written into a 100M-row table ----------
n_chunk, n_rep = 4 << 20, 25
x = torch.randn(n_chunk, 64, device=dev)
cb = codebooks(3, 256, 64)
table = torch.empty((3, n_chunk * n_rep), dtype=torch.int64, device=dev)


def tokenize_100m():
    for rep in range(n_rep):
        _, _, ids, _ = RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                           want_loss=False)
        table[:, rep * n_chunk:(rep + 1) * n_chunk].copy_(ids)


ms = timed(tokenize_100m, 2, warm=1)
out["tokenize_100M_items_d64_3x256_chunked"] = {
    "items": n_chunk * n_rep, "ms": ms, "items_per_s": n_chunk * n_rep / (ms * 1e-3),
    "note": "quantise-only, 4M-item resident chunk cycled 25x, [3, 100M] int64 id table written"}
del x, table
# ---- C4-like, host buffers: pinned latents -> ids through the C-ABI host entry (chunked 3-stream pipeline) ----
import time
n = 4 << 20
xh = torch.randn(n, 32).pin_memory()
cb = codebooks(3, 256, 32)
idh = torch.empty((3, n), dtype=torch.int64).pin_memory()


def host_call():
    rc = R.lib().rq_semantic_ids_host(xh.data_ptr(), cb.data_ptr(), n, 32, 256, 3, 1, idh.data_ptr(), None, 0, None)
    assert rc == 0


host_call()
host_call()
torch.cuda.synchronize()
times = []
for _ in range(7):
    t0 = time.perf_counter()
    host_call()
    times.append(time.perf_counter() - t0)
dt = sorted(times)[len(times) // 2]                     # median of 7 calls
out["tokenize_host_buffers_d32_3x256"] = {
    "items": n, "ms": dt * 1e3, "items_per_s": n / dt, "h2d_GBps": n * 128 / dt / 1e9, "d2h_GBps": n * 24 / dt / 1e9,
    "note": "wall clock around rq_semantic_ids_host (median of 7), pinned host buffers, copies inside",
    "ms_min_max": [min(times) * 1e3, max(times) * 1e3]}
del xh, idh
# ---- C4-like, module level: SemanticIdTokenizer.precompute_corpus_ids over 1M pinned 768-d items (ml32m dims) ------
from rq_vae_recommender_b200.data.schemas import SeqBatch


class _Corpus(torch.utils.data.Dataset):
    def __init__(self, n, dim):
        self.x = torch.randn(n, dim).pin_memory()

    def __len__(self):
        return self.x.shape[0]

    def __getitem__(self, idx):
        lo, hi = idx[0], idx[-1] + 1                      # the tokenizer asks for contiguous ranges
        ids = torch.arange(lo, hi)
        m1 = -torch.ones(hi - lo, dtype=torch.int64)
        return SeqBatch(user_ids=m1, ids=ids, ids_fut=m1, x=self.x[lo:hi], x_fut=m1, seq_mask=torch.ones(hi - lo, dtype=torch.bool))


n_items = 1 << 20
corpus = _Corpus(n_items, 768)
tok = R.SemanticIdTokenizer(input_dim=768, output_dim=64, hidden_dims=[512, 256, 128], codebook_size=256, n_layers=3,
                            n_cat_feats=0).to(dev)
with torch.no_grad():
    lat = tok.rq_vae.encode(corpus.x[:65536].to(dev))
    for l, layer in enumerate(tok.rq_vae.layers):        # trained-looking codebooks: residual rows of the data
        layer.embedding.weight.copy_(lat[l * 256:(l + 1) * 256] * (0.6 ** l))
        layer.kmeans_initted = True
tok.precompute_corpus_ids(corpus)
torch.cuda.synchronize()
t0 = time.perf_counter()
table = tok.precompute_corpus_ids(corpus)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
out["corpus_tokenisation_module_api_1M_x_768_to_d64_3x256"] = {
    "items": n_items, "ms": dt * 1e3, "items_per_s": n_items / dt, "h2d_GBps": n_items * 768 * 4 / dt / 1e9,
    "table_shape": list(table.shape),
    "note": "pinned 768-d inputs -> encoder MLP (torch) -> fused quantiser -> dedup table; wall clock, H2D inside"}
del corpus, tok, table
# ---- C3-like: one k-means level-iteration ------------------------------------------------------------------
n = 10_000_000
x = torch.randn(n, 32, device=dev)
c = x[torch.randperm(n, device=dev)[:256]].clone()
assign = torch.empty(n, dtype=torch.int64, device=dev)
RF.kmeans_assign_accumulate(x, c, assign)


def km_iter():
    a, sums, counts, nch = RF.kmeans_assign_accumulate(x, c, assign)
    RF.kmeans_update(sums, counts, c)


ms = timed(km_iter, 5)
out["kmeans_level_iteration_10M_d32_k256"] = {
    "points": n, "ms": ms, "points_per_s": n / (ms * 1e-3), "algorithmic_GBps": n * 128 / (ms * 1e-3) / 1e9,
    "frac_of_hbm_roofline": (n * 128 / (ms * 1e-3) / 1e9) / HBM}
import numpy as np
np.random.seed(0)
R.Kmeans(k=256, max_iters=3).run(x[:100000])


def km_run(iters):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    R.Kmeans(k=256, max_iters=iters).run(x)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3


t5, t25 = km_run(5), km_run(105)
out["kmeans_level_iteration_10M_d32_k256"]["Kmeans_run_ms_per_iter_incl_host"] = (t25 - t5) / 100
out["kmeans_level_iteration_10M_d32_k256"]["Kmeans_run_fixed_ms"] = t5 - 5 * (t25 - t5) / 100   # np.random.choice(10M, 256) etc.
ms_assign = timed(lambda: RF.quantize_forward(x, c[None], RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                              want_loss=False), 5)
out["kmeans_level_iteration_10M_d32_k256"]["assign_only_ms"] = ms_assign
ms_simt = timed(lambda: RF.kmeans_assign_accumulate(x, c, assign, impl=RF.IMPL_SIMT), 2, warm=1)
out["kmeans_level_iteration_10M_d32_k256"]["simt_family_ms"] = ms_simt
del x, assign
# ---- C5-like: 4 x 1024, d = 64 training step (rotation: chunked tcgen05 forward; Gumbel: SIMT family) ---------------
D, K, L = 64, 1024, 4
cb = codebooks(L, K, D)
for B in (16384, 65536):
    x = torch.randn(B, D, device=dev)
    g_emb = torch.randn(L, B, D, device=dev) * 0.01
    g_loss = torch.full((B,), 1.0 / B, device=dev)
    for name, mode in (("rotation", RF.MODE_ROTATION_TRICK), ("gumbel", RF.MODE_GUMBEL_SOFTMAX)):
        emb, res, ids, loss = RF.quantize_forward(x, cb, mode, 0.25, 0.2, seed=1, offset=2)
        f = timed(lambda: RF.quantize_forward(x, cb, mode, 0.25, 0.2, seed=1, offset=2), 3, warm=1)
        b = timed(lambda: RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, mode, 0.25, 0.2, seed=1, offset=2), 3, warm=1)
        out[f"train_step_{name}_4x1024_d64_B{B}"] = {"items": B, "fwd_ms": f, "bwd_ms": b, "ms": f + b,
                                                     "items_per_s": B / ((f + b) * 1e-3)}
print(json.dumps(out, indent=1))
