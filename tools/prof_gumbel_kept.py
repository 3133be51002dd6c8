"""Profiling driver: Gumbel-softmax forward + backward at the C5 shape with the softmax weights kept (run under ncu)."""
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
m = RF.MODE_GUMBEL_SOFTMAX
for i in range(2):
    w = RF.gumbel_weights_buffer(x, cb, m)
    emb, res, ids, loss = RF.quantize_forward(x, cb, m, 0.25, 0.2, seed=1, offset=2, gumbel_weights_out=w)
    RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, m, 0.25, 0.2, seed=1, offset=2, gumbel_weights=w)
torch.cuda.synchronize()
