"""Gumbel-softmax forward / backward, tensor-core path against the thread-per-row kernels, CUDA events.  Profiling only."""
import os, sys
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


for (B, D, K, L) in ((65536, 64, 1024, 4), (65536, 32, 256, 3), (16384, 64, 1024, 4)):
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.randn(B, D, device=dev, generator=g)
    cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
    g_emb = torch.randn(L, B, D, device=dev, generator=g) * 0.01
    g_loss = torch.full((B,), 1.0 / B, device=dev)
    M = RF.MODE_GUMBEL_SOFTMAX
    emb, res, ids, loss = RF.quantize_forward(x, cb, M, 0.25, 0.2, seed=1, offset=2)
    out = []
    w = RF.gumbel_weights_buffer(x, cb, M, impl=RF.IMPL_TCGEN05)
    RF.quantize_forward(x, cb, M, 0.25, 0.2, seed=1, offset=2, impl=RF.IMPL_TCGEN05, gumbel_weights_out=w)
    for name, impl, tpr, ww in (("tensor cores, weights kept", RF.IMPL_TCGEN05, False, w), ("tensor cores, recompute", RF.IMPL_TCGEN05, False, None),
                                ("thread per row", RF.IMPL_SIMT, True, None)):
        f = timed(lambda: RF.quantize_forward(x, cb, M, 0.25, 0.2, seed=1, offset=2, impl=impl, gumbel_weights_out=ww))
        b = timed(lambda: RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, M, 0.25, 0.2, seed=1, offset=2,
                                               thread_per_row=tpr, gumbel_weights=ww))
        out.append(f"{name}: fwd {f:.3f} + bwd {b:.3f} = {f + b:.3f} ms ({B / (f + b) / 1e3:.1f} M items/s)")
    print(f"B={B} D={D} K={K} L={L}  " + " | ".join(out))
