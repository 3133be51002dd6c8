"""Fuzz the near-tie / rescoring path of the tcgen05 family against SIMT and the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import rq_oracle as O
from rq_vae_recommender_b200 import functional as RF, _lib
from parity_utils import check_ids, synthetic_codebooks
DEV = "cuda:0"
bad = 0
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 20):
    for scale in (1e-3, 1.0, 50.0):
        g = torch.Generator().manual_seed(seed)
        B, K, L = 6000, 256, 3
        x = torch.randn(B, 32, generator=g) * scale
        cbs = synthetic_codebooks(torch.randn(2048, 32, generator=g) * scale, L, K, seed=seed + 1)
        cbs[0][17] = cbs[0][5]
        x[:200] = 0.5 * (cbs[0][torch.arange(200) % K] + cbs[0][(torch.arange(200) + 1) % K]) + 1e-7 * scale * torch.randn(200, 32, generator=g)
        x[200:260] = cbs[0][5] + 1e-4 * scale * torch.randn(60, 32, generator=g)
        cd = torch.stack(cbs).to(DEV)
        _, _, ids, _ = RF.quantize_forward(x.to(DEV), cd, RF.MODE_EVAL, impl=RF.IMPL_TCGEN05)
        _, _, ids_s, _ = RF.quantize_forward(x.to(DEV), cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
        nd = int((ids != ids_s).sum())
        msg = ""
        try:
            ref = O.rq_forward(x, cbs, O.MODE_EVAL)
            check_ids(ids_s, ref, cbs, "simt")
        except AssertionError as e:
            msg = str(e)[:300]
        if nd or msg or _lib.lib().rq_debug_tc_timeout_flag():
            bad += 1
            rows = torch.nonzero((ids != ids_s).any(0)).flatten()[:5].tolist()
            print(f"seed {seed} scale {scale}: tc!=simt {nd} rows {rows} flag {_lib.lib().rq_debug_tc_timeout_flag()} oracle: {msg}")
print("bad", bad)
