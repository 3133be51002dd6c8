"""Times one k-means level-iteration (assign + accumulate + update) and the assignment alone at the C3 shape
(10M x 32, K = 256) with CUDA events.  Profiling only.  usage: python tools/time_kmeans.py [B] [D] [K]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF, _lib
if os.environ.get("RQ_AB_LIB"):
    _lib.LIB_PATH = os.environ["RQ_AB_LIB"]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 32
K = int(sys.argv[3]) if len(sys.argv) > 3 else 256
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(B, D, device=dev, generator=g)
c = x[torch.randperm(B, device=dev, generator=g)[:K]].clone()
assign = torch.zeros(B, dtype=torch.int64, device=dev)


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def it():
    a, sums, counts, nch = RF.kmeans_assign_accumulate(x, c, assign)
    RF.kmeans_update(sums, counts, c)


ms = timed(it)
ms_assign = timed(lambda: RF.quantize_forward(x, c[None], RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                              want_loss=False))
hbm = B * D * 4 / 1e9
print(f"{os.environ.get('RQ_AB_LIB', 'default')}: level-iteration {ms:.3f} ms = {B / ms / 1e6:.2f} G points/s "
      f"({hbm / ms * 1e3:.0f} GB/s algorithmic); assignment alone {ms_assign:.3f} ms")
# sums / counts against torch
a, sums, counts, _ = RF.kmeans_assign_accumulate(x, c)
want = torch.zeros(K, D, device=dev, dtype=torch.float64).index_add_(0, a, x.double())
err = float((sums.double() - want).abs().max() / want.abs().max())
print(f"sums rel err vs float64 index_add: {err:.2e}; counts ok: {bool(torch.equal(counts, torch.bincount(a, minlength=K)))}")
