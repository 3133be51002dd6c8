"""Host-side cost of each segment of bench.py's e2e step (perf_counter between segments, no syncs except the last)."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import rq_vae_recommender_b200 as R

dev = torch.device("cuda:0")
N = 8
xs_h, cbs, g_embs, g_loss = bench.make_inputs(torch, dev, N, seed=1)
cbs = cbs.to(dev); g_embs = [g.to(dev) for g in g_embs]; g_loss = g_loss.to(dev)
B, D = xs_h[0].shape
L, K = cbs.shape[0], cbs.shape[1]
model = R.RqVae(input_dim=D, embed_dim=D, hidden_dims=[], codebook_size=K, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=L, commitment_weight=0.25,
                n_cat_features=0).to(dev)
with torch.no_grad():
    for l in range(L):
        model.layers[l].embedding.weight.copy_(cbs[l])
model.train()
xs_pin = [t.pin_memory() for t in xs_h]
loss_pin = torch.empty(B, dtype=torch.float32).pin_memory()
ids_pin = torch.empty((L, B), dtype=torch.int64).pin_memory()
feed = R.data.utils.DevicePrefetcher(R.data.utils.cycle(xs_pin), dev)
gb = [g.permute(1, 2, 0) for g in g_embs]
acc = {}


def mark(name, t0):
    t = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t - t0)
    return t


def step(i):
    s = i % N
    t = time.perf_counter()
    lat = next(feed).requires_grad_(True); t = mark("feed", t)
    out = model.quantize_latents(lat, 0.2); t = mark("forward", t)
    model.zero_grad(set_to_none=True); t = mark("zero_grad", t)
    torch.autograd.backward([out.embeddings, out.quantize_loss], [gb[s], g_loss]); t = mark("backward", t)
    loss_pin.copy_(out.quantize_loss.detach(), non_blocking=True)
    ids_pin.copy_(out.sem_ids.detach().t(), non_blocking=True); t = mark("d2h issue", t)
    torch.cuda.current_stream().synchronize(); t = mark("sync", t)


for i in range(5):
    step(i)
acc.clear()
n = 50
t0 = time.perf_counter()
for i in range(n):
    step(i)
tot = time.perf_counter() - t0
print(f"step {tot / n * 1e6:.0f} us")
for k, v in acc.items():
    print(f"  {k:12s} {v / n * 1e6:7.1f} us")
