import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if os.environ.get("CHILD"):
    import torch
    from rq_vae_recommender_b200 import _lib
    _lib.LIB_PATH = os.environ["LIBP"]
    from rq_vae_recommender_b200 import functional as RF
    dev = torch.device("cuda:0")
    for B in (16384, 65536):
        D, K, L = 64, 1024, 4
        g = torch.Generator(device=dev).manual_seed(0)
        x = torch.randn(B, D, device=dev, generator=g)
        cb = torch.stack([torch.randn(K, D, device=dev, generator=g) * (0.6 ** l) for l in range(L)])
        g_emb = torch.randn(L, B, D, device=dev, generator=g) * 0.01
        g_loss = torch.full((B,), 1.0 / B, device=dev)
        mode = RF.MODE_GUMBEL_SOFTMAX
        emb, res, ids, loss = RF.quantize_forward(x, cb, mode, 0.25, 0.2, seed=1, offset=2)
        f = lambda: RF.quantize_backward(g_emb, None, g_loss, res, ids, cb, mode, 0.25, 0.2, seed=1, offset=2)
        f(); f()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(4): f()
        b.record(); torch.cuda.synchronize()
        print(os.path.basename(os.environ["LIBP"]), "B", B, "bwd ms %.3f" % (a.elapsed_time(b) / 4), flush=True)
else:
    for v in sys.argv[1:]:
        subprocess.run([sys.executable, __file__], env=dict(os.environ, CHILD="1", LIBP=os.path.abspath(v)))
