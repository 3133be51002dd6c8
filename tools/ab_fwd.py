"""A/B timing of the forward kernel at the bench shape for alternative builds of the library.

    python tools/ab_fwd.py [path/to/librqengine_variant.so ...]

Each variant runs in its own subprocess (RQ_AB_LIB overrides the library path), graph replays over a ring of 8
batches like bench.py; prints us per launch for the full forward and the ids-only launch.  Profiling only."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def child():
    import torch
    from rq_vae_recommender_b200 import _lib
    if os.environ.get("RQ_AB_LIB"):
        _lib.LIB_PATH = os.environ["RQ_AB_LIB"]
    import bench
    from rq_vae_recommender_b200 import functional as RF
    dev = torch.device("cuda:0")
    N = 8
    xs, cbs, g_emb, g_loss = bench.make_inputs(torch, dev, N, seed=1)
    xs = [t.to(dev) for t in xs]; cbs = cbs.to(dev)
    mode = RF.MODE_ROTATION_TRICK
    out = {}
    g_emb = [t.to(dev) for t in g_emb]; g_loss = g_loss.to(dev)
    fo = [RF.quantize_forward(xs[s], cbs, mode, 0.25, 0.2) for s in range(N)]
    for name, fn in (("bwd", lambda s: RF.quantize_backward(g_emb[s], None, g_loss, fo[s][1], fo[s][2], cbs, mode, 0.25, 0.2)),
                     ("fwd", lambda s: RF.quantize_forward(xs[s], cbs, mode, 0.25, 0.2)),
                     ("ids", lambda s: RF.quantize_forward(xs[s], cbs, RF.MODE_EVAL, want_embeddings=False,
                                                           want_residuals=False, want_loss=False))):
        for s in range(N):
            fn(s)
        torch.cuda.synchronize()
        graphs, keep = [], []
        for s in range(N):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                keep.append(fn(s))
            graphs.append(g)
        for i in range(16):
            graphs[i % N].replay()
        best = 1e9
        for rep in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(40):
                graphs[i % N].replay()
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 40 * 1e3)
        out[name] = best
        if name == "fwd":
            # the same launches back to back inside ONE graph (40 kernel nodes): per-launch time without the
            # graph-launch gap that a one-kernel graph pays
            gg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gg):
                for i in range(40):
                    keep.append(fn(i % N))
            gg.replay()
            bestg = 1e9
            for rep in range(5):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                gg.replay()
                e1.record()
                torch.cuda.synchronize()
                bestg = min(bestg, e0.elapsed_time(e1) / 40 * 1e3)
            out["fwd_in_graph"] = bestg
    ref = keep[0]
    print(f"{os.environ.get('RQ_AB_LIB', 'default'):50s} fwd {out['fwd']:.2f} us (40 in one graph: {out['fwd_in_graph']:.2f})  ids-only {out['ids']:.2f} us  bwd {out['bwd']:.2f} us", flush=True)


if __name__ == "__main__":
    if os.environ.get("RQ_AB_CHILD"):
        child()
    else:
        variants = sys.argv[1:] or [""]
        for v in variants:
            env = dict(os.environ, RQ_AB_CHILD="1")
            if v:
                env["RQ_AB_LIB"] = os.path.abspath(v)
            subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)
