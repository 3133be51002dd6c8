"""Large-launch throughput of the d=32 forward (many tiles per warpgroup): ids-only and full outputs, 4M rows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF, _lib
if os.environ.get("RQ_AB_LIB"):
    _lib.LIB_PATH = os.environ["RQ_AB_LIB"]
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
n = 4 << 20
x = torch.randn(n, 32, device=dev)
cb = torch.stack([torch.randn(256, 32, generator=g) * (0.6 ** l) for l in range(3)]).to(dev)


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


ms = timed(lambda: RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False))
print(f"ids-only 4M x 32, 3x256: {ms:.3f} ms = {n / ms / 1e6:.2f} G items/s")
for mode, name in ((RF.MODE_ROTATION_TRICK, "rotation"), (RF.MODE_EVAL, "eval")):
    ms = timed(lambda: RF.quantize_forward(x, cb, mode, 0.25, 0.2))
    print(f"full forward ({name}) 4M x 32: {ms:.3f} ms = {n / ms / 1e6:.2f} G items/s = {n * 924 / ms / 1e6:.0f} GB/s algorithmic")
