"""Profiling driver: one Gumbel-softmax forward + backward at the C5 shape (run under ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF
dev = torch.device("cuda:0")
B, D, K, L = 65536, 64, 1024, 4
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(B, D, device=dev, generator=g)
cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
g_emb = torch.randn(L, B, D, device=dev, generator=g) * 0.01
g_loss = torch.full((B,), 1.0 / B, device=dev)
for i in range(2):
    emb, res, ids, loss = RF.quantize_forward(x, cb, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, seed=1, offset=2)
    RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, seed=1, offset=2)
torch.cuda.synchronize()
