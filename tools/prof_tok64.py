"""Profiling driver: ids-only tokenisation launches of a 1M-item d = 64, 3 x 256 chunk (run under ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
D, K, L = 64, 256, 3
cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
x = torch.randn(1_000_000, D, device=dev, generator=g)
for i in range(3):
    RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False)
torch.cuda.synchronize()
