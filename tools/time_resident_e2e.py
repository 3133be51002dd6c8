"""Device-resident 768-d -> semantic ids (SURVEY 8d rows "end-to-end 768 -> ids"): the encoder MLP on cuBLAS through
torch (fp32 'highest' and TF32 'high', the reference's setting, modules/rqvae.py:19) + the fused quantiser, 1M resident
items in 262,144-item chunks.  Profiling only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rq_vae_recommender_b200 as R

dev = torch.device("cuda:0")
N, CH = 1 << 20, 1 << 18
for d, hidden in ((64, [512, 256, 128]), (32, [512, 256, 128])):
    torch.manual_seed(0)
    model = R.RqVae(input_dim=768, embed_dim=d, hidden_dims=hidden, codebook_size=256, codebook_kmeans_init=False,
                    codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0).to(dev).eval()
    with torch.no_grad():
        for l, layer in enumerate(model.layers):
            layer.embedding.weight.copy_(torch.randn_like(layer.embedding.weight) * (0.6 ** l))
    x = torch.nn.functional.normalize(torch.randn(N, 768, device=dev), dim=-1)
    for prec, eng in (("highest", "torch"), ("high", "torch"), ("high", "engine")):
        torch.set_float32_matmul_precision(prec)
        os.environ["RQ_ENCODER_TORCH"] = "0" if eng == "engine" else "1"

        def run():
            with torch.no_grad():
                for s in range(0, N, CH):
                    model.get_semantic_ids(x[s:s + CH])

        def enc():
            with torch.no_grad():
                for s in range(0, N, CH):
                    model.encode(x[s:s + CH])
        out = []
        for fn in (run, enc):
            fn(); fn()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            for _ in range(3):
                fn()
            b.record()
            torch.cuda.synchronize()
            out.append(a.elapsed_time(b) / 3)
        print(f"d={d} matmul precision {prec:8s} encoder on {eng:6s}: 768 -> ids {out[0]:.2f} ms per 1M items = {N / out[0] / 1e3:.1f} M items/s "
              f"(encoder alone {out[1]:.2f} ms, quantiser + glue {out[0] - out[1]:.2f} ms)")
