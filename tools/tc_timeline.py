"""Prints the clock64 phase timeline of CTA 0 of the tcgen05 forward kernel (rq_debug_tc_set_timeline)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from rq_vae_recommender_b200 import functional as RF, _lib
if os.environ.get("RQ_AB_LIB"):                      # an -DRQ_EXPERIMENTS build of the library (build.py --experiments)
    _lib.LIB_PATH = os.environ["RQ_AB_LIB"]

dev = torch.device("cuda:0")
xs, cbs, g_emb, g_loss = bench.make_inputs(torch, dev, 2, seed=1)
xs = [t.to(dev) for t in xs]; cbs = cbs.to(dev)
mode = RF.MODE_ROTATION_TRICK
for i in range(3):
    RF.quantize_forward(xs[i % 2], cbs, mode, 0.25, 0.2)
buf = torch.zeros(32, 64, dtype=torch.int64, device=dev)
_lib.lib().rq_debug_tc_set_timeline(buf.data_ptr())
if os.environ.get("TL_WARM_X"):
    s = xs[1].sum().item()          # x tile source now L2-resident, GPU idle before the launch
if os.environ.get("TL_SYNC"):
    torch.cuda.synchronize()
RF.quantize_forward(xs[1], cbs, mode, 0.25, 0.2)
torch.cuda.synchronize()
_lib.lib().rq_debug_tc_set_timeline(None)
t = buf.cpu()
t0 = int(t[t > 0].min())
names = ["start", "init", "conv", "stagein"] + [f"L{l}:{n}" for l in range(3) for n in
                                              ("issued", "gathered", "accrdy", "scanned", "id", "est", "Awr", "pwait")]
print("x landed:", [int(t[w, 60]) - t0 for w in range(16)])
print("warp " + " ".join(f"{n:>9s}" for n in names))
for w in range(32):
    if int(t[w, 0]) == 0:
        continue
    print(f"{w:4d} " + " ".join(f"{(int(v) - t0) if int(v) else -1:9d}" for v in t[w, :len(names)]))
