"""Profiling driver: a few launches of the fused quantiser at the C2 shape (run under ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from rq_vae_recommender_b200 import functional as RF

mode = {"rot": RF.MODE_ROTATION_TRICK, "eval": RF.MODE_EVAL}[sys.argv[1] if len(sys.argv) > 1 else "rot"]
what = sys.argv[2] if len(sys.argv) > 2 else "fwd"
dev = torch.device("cuda:0")
xs, cbs, g_emb, g_loss = bench.make_inputs(torch, dev, 2, seed=1)
xs = [t.to(dev) for t in xs]; cbs = cbs.to(dev); g_emb = [t.to(dev) for t in g_emb]; g_loss = g_loss.to(dev)
for i in range(6):
    if what == "ids":
        RF.quantize_forward(xs[i % 2], cbs, RF.MODE_EVAL, want_embeddings=False, want_residuals=False, want_loss=False)
    else:
        emb, res, ids, loss = RF.quantize_forward(xs[i % 2], cbs, mode, 0.25, 0.2)
        if what == "fwdbwd":
            RF.quantize_backward(g_emb[i % 2], None, g_loss, res, ids, cbs, mode, 0.25, 0.2)
torch.cuda.synchronize()
print("ok")
