"""Timing experiments for the backward kernels at the C2 shape: RQ_BWD_DBG bits switch phases off
(1 no REDG, 2 no g_x store, 4 no math, 8 no loads); results are wrong with any bit set — timing only."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from rq_vae_recommender_b200 import functional as RF

dev = torch.device("cuda:0")
N = 8
xs, cbs, g_emb, g_loss = bench.make_inputs(torch, dev, N, seed=1)
xs = [t.to(dev) for t in xs]; cbs = cbs.to(dev); g_emb = [t.to(dev) for t in g_emb]; g_loss = g_loss.to(dev)
mode = RF.MODE_ROTATION_TRICK
fw = [RF.quantize_forward(xs[i], cbs, mode, 0.25, 0.2) for i in range(N)]


def run(label, env):
    for k in ("RQ_BWD_DBG", "RQ_BWD_NO_ROW3", "RQ_BWD_NO_ROW2", "RQ_BWD_NO_ROW"):
        os.environ.pop(k, None)
    os.environ.update(env)
    def go(i):
        emb, res, ids, loss = fw[i % N]
        RF.quantize_backward(g_emb[i % N], None, g_loss, res, ids, cbs, mode, 0.25, 0.2)
    for i in range(3):
        go(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(N):
            go(i)
    g.replay()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for i in range(5):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    print(f"{label:40s} {a.elapsed_time(b) / (5 * N) * 1e3:8.1f} us per backward (graph replay, memset + kernel)")


run("row3 (2 lanes/row, double-buffered)", {})
R2 = {"RQ_BWD_NO_ROW3": "1"}
run("row2 full", dict(R2))
run("row2 no REDG", dict(R2, RQ_BWD_DBG="1"))
run("row2 no g_x store", dict(R2, RQ_BWD_DBG="2"))
run("row2 no math", dict(R2, RQ_BWD_DBG="4"))
run("row2 no loads", dict(R2, RQ_BWD_DBG="8"))
run("row2 nothing", dict(R2, RQ_BWD_DBG="15"))
run("row (ring) full", {"RQ_BWD_NO_ROW3": "1", "RQ_BWD_NO_ROW2": "1"})
