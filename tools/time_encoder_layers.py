"""Per-layer time of the encoder-layer kernel against torch (cuBLAS TF32 matmul + SiLU) at the reference encoder's
shapes, 1M rows.  Profiling only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")
M = int(os.environ.get("ROWS", 1 << 20))
torch.set_float32_matmul_precision("high")


def timed(fn, n=5):
    fn(); fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


ONLY = os.environ.get("LAYER")
for n_in, n_out, act in ((768, 512, 1), (768, 512, 0), (768, 256, 0), (768, 128, 0), (1536, 512, 0), (512, 256, 1), (256, 128, 1), (128, 64, 0), (128, 32, 0), (128, 32, 2)):
    if ONLY and ONLY != f"{n_in}x{n_out}":
        continue
    x = torch.randn(M, n_in, device=dev)
    w = torch.randn(n_out, n_in, device=dev) / n_in ** 0.5
    y = torch.empty(M, n_out, device=dev)
    te = timed(lambda: RF.encoder_layer(x, w, act | (RF.ACT_CTA_PAIR if os.environ.get('ENC_PAIR') else 0), out=y))
    if act == 1:
        tt = timed(lambda: torch.nn.functional.silu(x @ w.t()))
    elif act == 2:
        tt = timed(lambda: torch.nn.functional.normalize(x @ w.t(), dim=-1))
    else:
        tt = timed(lambda: x @ w.t())
    tg = timed(lambda: torch.matmul(x, w.t(), out=y))
    fl = 2.0 * M * n_in * n_out
    by = 4.0 * M * (n_in + n_out)
    print(f"{n_in:4d}->{n_out:3d} act {act}: engine {te:.3f} ms ({fl / te / 1e9:.0f} TF/s, {by / te / 1e6:.0f} GB/s)   "
          f"torch {tt:.3f} ms (matmul alone {tg:.3f} ms)")
