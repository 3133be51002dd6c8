"""Where does a torch.compile(mode="reduce-overhead") step of the quantiser module spend its time?  (profiling only)
    TORCH_LOGS=perf_hints,cudagraphs python tools/compile_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")
B, D, K, L = 65536, 32, 256, 3
g = torch.Generator().manual_seed(0)
model = R.RqVae(input_dim=D, embed_dim=D, hidden_dims=[], codebook_size=K, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=L, commitment_weight=0.25,
                n_cat_features=0).to(dev)
model.train()


class QuantizeLatents(torch.nn.Module):
    def __init__(self, rqvae):
        super().__init__()
        self.rqvae = rqvae

    def forward(self, lat):
        out = self.rqvae.quantize_latents(lat, 0.2)
        return out.embeddings, out.sem_ids, out.quantize_loss


q = QuantizeLatents(model)
cq = torch.compile(q, mode=os.environ.get("MODE", "reduce-overhead"))
params = [p for p in model.parameters() if p.requires_grad]
g_emb = (torch.randn(L, B, D, generator=g) * 0.01).to(dev).permute(1, 2, 0)
if os.environ.get("CONTIG"):
    g_emb = g_emb.contiguous()
g_loss = torch.full((B,), 1.0 / B, device=dev)
lats = [torch.randn(B, D, device=dev) for _ in range(2)]


def step(i):
    lat = lats[i % 2].detach().requires_grad_(True)
    for p in params:
        p.grad = None
    torch.compiler.cudagraph_mark_step_begin()
    emb, ids, loss = cq(lat)
    torch.autograd.backward([emb, loss], [g_emb, g_loss])
    return ids, loss


for i in range(6):
    step(i)
torch.cuda.synchronize()
n = 50
t0 = time.perf_counter()
for i in range(n):
    step(i)
t_host = time.perf_counter() - t0
torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"host {t_host / n * 1e6:.1f} us/step, wall {t_all / n * 1e6:.1f} us/step")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(4):
        step(i)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
