"""Per-CTA wall-clock stamps (%globaltimer at entry / exit, %smid) of the tcgen05 forward kernel, next to the
CUDA-event duration of the same launches.  Needs an -DRQ_EXPERIMENTS build (RQ_AB_LIB=path).  Profiling only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from rq_vae_recommender_b200 import functional as RF, _lib
if os.environ.get("RQ_AB_LIB"):
    _lib.LIB_PATH = os.environ["RQ_AB_LIB"]

dev = torch.device("cuda:0")
N = 8
xs, cbs, g_emb, g_loss = bench.make_inputs(torch, dev, N, seed=1)
xs = [t.to(dev) for t in xs]; cbs = cbs.to(dev)
mode = RF.MODE_ROTATION_TRICK
for i in range(5):
    RF.quantize_forward(xs[i % N], cbs, mode, 0.25, 0.2)
buf = torch.zeros(4096, dtype=torch.int64, device=dev)
_lib.lib().rq_debug_tc_set_timeline(buf.data_ptr())
torch.cuda.synchronize()
res = []
for rep in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    buf.zero_()
    torch.cuda.synchronize()
    nq = int(os.environ.get("TC_QUEUE", "1"))          # launches queued back to back; stamps are the last one's
    e0.record()
    for j in range(nq):
        RF.quantize_forward(xs[(rep + j) % N], cbs, mode, 0.25, 0.2)
    e1.record()
    torch.cuda.synchronize()
    t = buf.cpu()[2048:2048 + 4 * 148].view(148, 4)
    st, sm, en = t[:, 0], t[:, 1], t[:, 2]
    t0 = int(st.min())
    dur = (en - st).float() / 1e3
    print(f"rep {rep}: event {e0.elapsed_time(e1) * 1e3 / nq:.1f} us/launch | first start -> last end {(int(en.max()) - t0) / 1e3:.2f} us | "
          f"start spread {(int(st.max()) - t0) / 1e3:.2f} us | CTA dur min/med/max {dur.min():.2f}/{dur.median():.2f}/{dur.max():.2f} us")
    if rep == 5:
        order = torch.argsort(en)
        print("slowest CTAs (bid, smid, start_us, dur_us):", [(int(b), int(sm[b]), round((int(st[b]) - t0) / 1e3, 2), round(float(dur[b]), 2)) for b in order[-8:]])
        print("fastest CTAs:", [(int(b), int(sm[b]), round((int(st[b]) - t0) / 1e3, 2), round(float(dur[b]), 2)) for b in order[:8]])
        four = torch.tensor([(512 - 1 - b) // 148 + 1 for b in range(148)])
        for n in (3, 4):
            m = four == n
            print(f"CTAs with {n} tiles: n={int(m.sum())}, dur mean {dur[m].mean():.2f} max {dur[m].max():.2f}")
_lib.lib().rq_debug_tc_set_timeline(None)
