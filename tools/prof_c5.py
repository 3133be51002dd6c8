import os, sys
sys.path.insert(0, "/root/repo")
import torch
from rq_vae_recommender_b200 import functional as RF
dev = torch.device("cuda:0")
B, D, K, L = 65536, 64, 1024, 4
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(B, D, device=dev, generator=g)
cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
for i in range(3):
    RF.quantize_forward(x, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2)
torch.cuda.synchronize()
