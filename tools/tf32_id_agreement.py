"""How many items keep their semantic ids when the encoder runs in TF32 (engine kernel, or torch's cuBLAS TF32) instead of
fp32 ('highest')?  Reference encoder 768-512-256-128-32, 3 x 256 codes (k-means-like codebooks from the latents), 262,144
items.  Profiling / documentation only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rq_vae_recommender_b200 as R

dev = torch.device("cuda:0")
torch.manual_seed(0)
N = 1 << 18
model = R.RqVae(input_dim=768, embed_dim=32, hidden_dims=[512, 256, 128], codebook_size=256, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0).to(dev).eval()
x = torch.nn.functional.normalize(torch.randn(N, 768, device=dev), dim=-1)
with torch.no_grad():
    torch.set_float32_matmul_precision("highest")
    z = model.encode(x)
    r = z.clone()
    for l, layer in enumerate(model.layers):          # codebooks = samples of the residuals, like a k-means init
        pick = torch.randperm(N, device=dev)[:256]
        layer.embedding.weight.copy_(r[pick])
        d = torch.cdist(r, layer.embedding.weight)
        r = r - layer.embedding.weight[d.argmin(1)]
    ids_fp32 = model.get_semantic_ids(x).sem_ids
    res = {}
    for name, env in (("torch TF32", "1"), ("engine TF32", "0")):
        torch.set_float32_matmul_precision("high")
        os.environ["RQ_ENCODER_TORCH"] = env
        ids = model.get_semantic_ids(x).sem_ids
        same_all = (ids == ids_fp32).all(dim=-1).float().mean().item()
        per_level = (ids == ids_fp32).float().mean(dim=0).tolist()
        zz = model.encode(x)
        rel = ((zz - z).norm() / z.norm()).item()
        print(f"{name:12s}: items with all 3 ids equal to the fp32 encoder's: {100 * same_all:.2f} %  per level "
              f"{[round(100 * p, 2) for p in per_level]}  latent rel. error {rel:.2e}")
