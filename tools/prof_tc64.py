"""Profiling driver: ids-only tokenisation at the ml32m shape (d=64, 3x256), 1M items (run under ncu)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")
n = 1 << 20
g = torch.Generator(device=dev).manual_seed(4)
x = torch.randn(n, 64, device=dev, generator=g)
cb = torch.stack([torch.randn(256, 64, device=dev, generator=g) * (0.6 ** l) for l in range(3)])
for i in range(3):
    RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False)
torch.cuda.synchronize()
print("ok")
