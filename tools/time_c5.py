"""Forward / backward split at the C5 shape (4 x 1024, d = 64) for rotation trick and Gumbel-softmax."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


for B in (16384, 65536):
    D, K, L = 64, 1024, 4
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.randn(B, D, device=dev, generator=g)
    cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
    g_emb = torch.randn(L, B, D, device=dev, generator=g) * 0.01
    g_loss = torch.full((B,), 1.0 / B, device=dev)
    for name, mode in (("rotation", RF.MODE_ROTATION_TRICK), ("gumbel", RF.MODE_GUMBEL_SOFTMAX), ("eval ids-only", RF.MODE_EVAL)):
        if mode == RF.MODE_EVAL:
            f = timed(lambda: RF.quantize_forward(x, cb, mode, want_embeddings=False, want_residuals=False, want_loss=False))
            print(f"B={B} {name:14s} fwd {f:8.3f} ms")
            continue
        emb, res, ids, loss = RF.quantize_forward(x, cb, mode, 0.25, 0.2, seed=1, offset=2)
        f = timed(lambda: RF.quantize_forward(x, cb, mode, 0.25, 0.2, seed=1, offset=2))
        b = timed(lambda: RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, mode, 0.25, 0.2, seed=1, offset=2))
        print(f"B={B} {name:14s} fwd {f:8.3f} ms   bwd {b:8.3f} ms")
