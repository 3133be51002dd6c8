"""ids-only tokenisation of a 4M-item d = 64, 3 x 256 chunk (configs[3] quantiser part) and the full training
forward at 65,536 items, CUDA events.  Profiling only.  RQ_AB_LIB selects another build of the library."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF, _lib
if os.environ.get("RQ_AB_LIB"):
    _lib.LIB_PATH = os.environ["RQ_AB_LIB"]
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
D, K, L = 64, 256, 3
cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])


def timed(fn, n=10, warm=3):
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


x = torch.randn(4_000_000, D, device=dev, generator=g)
ms = timed(lambda: RF.quantize_forward(x, cb, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False))
x2 = x[:65536].contiguous()
ms2 = timed(lambda: RF.quantize_forward(x2, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2), n=20)
print(f"{os.environ.get('RQ_AB_LIB', 'default')}: ids-only 4M x 64: {ms:.3f} ms = {4e6 / ms / 1e6:.2f} G items/s; "
      f"rotation forward 65,536 x 64: {ms2 * 1e3:.1f} us")
