"""GPU parity: fused quantiser (through the module API and the C ABI) vs golden fixtures and the
CPU oracle.  Run on the B200 box with `-m gpu`."""
import numpy as np
import pytest
import torch

import rq_oracle as O
import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import functional as RF
from conftest import load_golden
from parity_utils import assert_close, check_ids, matching_rows, synthetic_codebooks

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MODES = {"eval": O.MODE_EVAL, "ste": O.MODE_STE, "rot": O.MODE_ROTATION_TRICK, "gumbel": O.MODE_GUMBEL_SOFTMAX}
IMPLS = [RF.IMPL_SIMT, RF.IMPL_AUTO]


def _as_oracle_out(z, mname):
    return O.RqOut(torch.from_numpy(z[f"{mname}_emb"]), torch.from_numpy(z[f"{mname}_res"]),
                   torch.from_numpy(z[f"{mname}_ids"]), torch.from_numpy(z[f"{mname}_loss"]))


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("name", ["rq_stack_d32_k256_l3.npz", "rq_stack_d64_k256_l3.npz", "rq_stack_d64_k1024_l4.npz"])
def test_golden_stack_forward_backward(name, impl):
    """Fixtures produced by the unmodified reference (RqVae.get_semantic_ids + autograd)."""
    z = load_golden(name)
    x = torch.from_numpy(z["x"]).to(DEV)
    cbs_cpu = torch.from_numpy(z["codebooks"])
    cbs = cbs_cpu.to(DEV)
    g_emb, g_loss = torch.from_numpy(z["g_emb"]).to(DEV), torch.from_numpy(z["g_loss"]).to(DEV)
    for mname, mode in MODES.items():
        if f"{mname}_ids" not in z:
            continue
        ref = _as_oracle_out(z, mname)
        gu = torch.from_numpy(z["gumbel_uniform"]).to(DEV) if mname == "gumbel" else None
        xin = x.clone().requires_grad_(True)
        cin = cbs.clone().requires_grad_(True)
        emb, res, ids, loss = RF.residual_quantize(xin, cin, mode, 0.25, 0.2, RF.DIST_L2, gu, impl=impl)
        n_mis, _ = check_ids(ids, ref, list(cbs_cpu), f"{name}/{mname}")
        ok = matching_rows(ids, ref)
        assert ok.float().mean() > 0.99
        assert_close(emb.detach().cpu()[:, ok], ref.embeddings[:, ok], f"{mname} emb")
        assert_close(res.detach().cpu()[:, ok], ref.residuals[:, ok], f"{mname} residuals")
        assert_close(loss.detach().cpu()[ok], ref.quantize_loss[ok], f"{mname} loss")
        if mname == "gumbel":
            continue   # backward covered in test_gpu_gumbel.py
        ((emb * g_emb).sum() + (loss * g_loss).sum()).backward()
        assert_close(xin.grad.cpu()[ok], torch.from_numpy(z[f"{mname}_gx"])[ok], f"{mname} grad x")
        if n_mis == 0:
            assert_close(cin.grad.cpu(), torch.from_numpy(z[f"{mname}_gcb"]), f"{mname} grad codebooks")


@pytest.mark.parametrize("name", ["quantize_level_d32_k256.npz", "quantize_level_d64_k1024.npz"])
def test_golden_single_level_module(name):
    """`Quantize.forward` itself, all four branches + COSINE, against the reference's outputs."""
    z = load_golden(name)
    x = torch.from_numpy(z["x"]).to(DEV)
    cb = torch.from_numpy(z["codebook"])
    D, K = cb.shape[1], cb.shape[0]
    g_emb, g_loss = torch.from_numpy(z["g_emb"]).to(DEV), torch.from_numpy(z["g_loss"]).to(DEV)
    R.GumbelNoise.source = "torch"
    try:
        for mname, fm, training in [("eval", R.QuantizeForwardMode.STE, False), ("ste", R.QuantizeForwardMode.STE, True),
                                    ("rot", R.QuantizeForwardMode.ROTATION_TRICK, True)]:
            layer = R.Quantize(D, K, do_kmeans_init=False, forward_mode=fm, commitment_weight=0.25).to(DEV)
            with torch.no_grad():
                layer.embedding.weight.copy_(cb)
            layer.train(training)
            xin = x.clone().requires_grad_(True)
            out = layer(xin, temperature=0.2)
            assert out.ids.dtype == torch.int64 and out.embeddings.shape == x.shape and out.loss.shape == (x.shape[0],)
            same = out.ids.cpu() == torch.from_numpy(z[f"{mname}_ids"])
            assert same.float().mean() > 0.99
            assert int(out.ids[5]) == 3          # duplicated codeword rows 3/7: first index wins
            assert_close(out.embeddings.detach().cpu()[same], torch.from_numpy(z[f"{mname}_emb"])[same], f"{mname} emb")
            assert_close(out.loss.detach().cpu()[same], torch.from_numpy(z[f"{mname}_loss"])[same], f"{mname} loss")
            ((out.embeddings * g_emb).sum() + (out.loss * g_loss).sum()).backward()
            assert_close(xin.grad.cpu()[same], torch.from_numpy(z[f"{mname}_gx"])[same], f"{mname} gx")
            if bool(same.all()):
                assert_close(layer.embedding.weight.grad.cpu(), torch.from_numpy(z[f"{mname}_gcb"]), f"{mname} gcb")
        layer = R.Quantize(D, K, do_kmeans_init=False, forward_mode=R.QuantizeForwardMode.STE,
                           distance_mode=R.QuantizeDistance.COSINE).to(DEV).eval()
        with torch.no_grad():
            layer.embedding.weight.copy_(cb)
        out = layer(x, 0.2)
        same = out.ids.cpu() == torch.from_numpy(z["cosine_ids"])
        assert same.float().mean() > 0.99
        assert_close(out.loss.cpu()[same], torch.from_numpy(z["cosine_loss"])[same], "cosine loss")
    finally:
        R.GumbelNoise.source = "philox"


def test_golden_norm_simvq_module_stack():
    """codebook_normalize (level 0) + sim_vq through RqVae, gradients to embedding AND out_proj."""
    z = load_golden("rq_stack_norm_simvq_d16_k32_l2.npz")
    D, K, L = 16, 32, 2
    x = torch.from_numpy(z["x"]).to(DEV)
    for mname, fm, training in [("eval", R.QuantizeForwardMode.STE, False), ("ste", R.QuantizeForwardMode.STE, True),
                                ("rot", R.QuantizeForwardMode.ROTATION_TRICK, True)]:
        m = R.RqVae(input_dim=D, embed_dim=D, hidden_dims=[], codebook_size=K, codebook_kmeans_init=False,
                    codebook_normalize=True, codebook_sim_vq=True, codebook_mode=fm, n_layers=L,
                    commitment_weight=0.25, n_cat_features=0).to(DEV)
        with torch.no_grad():
            m.encoder.mlp[0].weight.copy_(torch.eye(D))
            for l in range(L):
                m.layers[l].embedding.weight.copy_(torch.from_numpy(z["codebooks"][l]))
                m.layers[l].out_proj[0].weight.copy_(torch.from_numpy(z["sim_vq_weights"][l]))
        m.train(training)
        xin = x.clone().requires_grad_(True)
        out = m.get_semantic_ids(xin, gumbel_t=0.2)
        B = x.shape[0]
        assert out.embeddings.shape == (B, D, L) and out.embeddings.stride() == (D, 1, B * D)
        assert out.sem_ids.shape == (B, L) and out.sem_ids.stride() == (1, B)
        same = (out.sem_ids.cpu() == torch.from_numpy(z[f"{mname}_ids"]).T).all(dim=1)
        assert same.float().mean() > 0.97
        assert_close(out.embeddings.permute(2, 0, 1).detach().cpu()[:, same], torch.from_numpy(z[f"{mname}_emb"])[:, same], "emb")
        assert_close(out.quantize_loss.detach().cpu()[same], torch.from_numpy(z[f"{mname}_loss"])[same], "loss")
        obj = (out.embeddings.permute(2, 0, 1) * torch.from_numpy(z["g_emb"]).to(DEV)).sum() + \
              (out.quantize_loss * torch.from_numpy(z["g_loss"]).to(DEV)).sum()
        obj.backward()
        assert_close(xin.grad.cpu()[same], torch.from_numpy(z[f"{mname}_gx"])[same], f"{mname} gx", tol=2e-3)
        if bool(same.all()):
            gcb = torch.stack([m.layers[l].embedding.weight.grad for l in range(L)]).cpu()
            assert_close(gcb, torch.from_numpy(z[f"{mname}_gcb"]), f"{mname} gcb", tol=2e-3)


@pytest.mark.parametrize("impl", IMPLS)
@pytest.mark.parametrize("B,D,K,L", [(1, 32, 256, 3), (255, 32, 256, 3), (1000, 32, 250, 2), (777, 64, 256, 3),
                                      (513, 16, 7, 4), (300, 8, 16, 1), (20000, 32, 256, 3)])
def test_seeded_vs_oracle_shapes(B, D, K, L, impl):
    """Ragged batch sizes, K not a multiple of 4, several D; ids-only launch agrees with full."""
    g = torch.Generator().manual_seed(B + D + K)
    x = torch.randn(B, D, generator=g)
    cbs = synthetic_codebooks(torch.randn(max(B, 2 * K), D, generator=g), L, K, seed=1)
    for mode in (O.MODE_EVAL, O.MODE_ROTATION_TRICK, O.MODE_STE):
        ref = O.rq_forward(x, cbs, mode)
        emb, res, ids, loss = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), mode, impl=impl)
        check_ids(ids, ref, cbs, f"B{B} D{D} K{K} mode{mode}")
        ok = matching_rows(ids, ref)
        assert_close(emb.cpu()[:, ok], ref.embeddings[:, ok], "emb")
        assert_close(res.cpu()[:, ok], ref.residuals[:, ok], "res")
        assert_close(loss.cpu()[ok], ref.quantize_loss[ok], "loss")
    _, _, ids2, _ = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), O.MODE_EVAL, want_embeddings=False,
                                        want_residuals=False, want_loss=False, impl=impl)
    _, _, ids1, _ = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), O.MODE_EVAL, impl=impl)
    assert torch.equal(ids1, ids2)


def test_empty_batch_and_errors():
    cbs = torch.randn(3, 256, 32, device=DEV)
    emb, res, ids, loss = RF.quantize_forward(torch.zeros(0, 32, device=DEV), cbs)
    assert emb.shape == (3, 0, 32) and ids.shape == (3, 0) and loss.shape == (0,)
    with pytest.raises(AssertionError):
        RF.quantize_forward(torch.zeros(4, 16, device=DEV), cbs)
    with pytest.raises(R.RqEngineError, match="unsupported"):
        RF.quantize_forward(torch.zeros(4, 24, device=DEV), torch.randn(1, 8, 24, device=DEV))
    with pytest.raises(R.RqEngineError):
        RF.quantize_forward(torch.zeros(4, 32), cbs)


def test_full_size_properties_c2():
    """BASELINE config 2 shape (B = 65,536, d=32, 3x256, ROTATION_TRICK): size-independent
    invariants instead of an oracle run."""
    B, D, K, L = 65536, 32, 256, 3
    g = torch.Generator().manual_seed(0)
    x = torch.randn(B, D, generator=g)
    cbs = torch.stack(synthetic_codebooks(x[:20000], L, K, seed=2)).to(DEV)
    xd = x.to(DEV)
    for mode in (O.MODE_EVAL, O.MODE_ROTATION_TRICK):
        emb, res, ids, loss = RF.quantize_forward(xd, cbs, mode, beta=0.25)
        assert int(ids.min()) >= 0 and int(ids.max()) < K
        assert torch.equal(res[0], xd)
        for l in range(L - 1):                   # residual chain, rqvae.py:129 — exact
            assert torch.equal(res[l + 1], res[l] - emb[l])
        e = torch.stack([cbs[l][ids[l]] for l in range(L)])
        want = sum(1.25 * ((res[l] - e[l]) ** 2).sum(-1) for l in range(L))
        assert_close(loss, want, "loss identity", tol=1e-5)
        if mode == O.MODE_EVAL:
            assert torch.equal(emb, e)           # eval branch returns the codeword bit-exactly
        else:
            assert_close(emb, e, "rotation value ~ codeword", tol=1e-4)
        # chosen code is (numerically) the nearest: no other code is closer by more than fp32 noise
        for l in range(L):
            d = torch.cdist(res[l][:4096].double(), cbs[l].double()) ** 2
            dmin = d.min(1).values
            dsel = d.gather(1, ids[l][:4096, None]).squeeze(1)
            scale = (res[l][:4096].double() ** 2).sum(1) + (e[l][:4096].double() ** 2).sum(1)
            assert bool(((dsel - dmin) <= 1e-6 * scale + 1e-12).all())
        # determinism / idempotence of the launch
        emb2, res2, ids2, loss2 = RF.quantize_forward(xd, cbs, mode, beta=0.25)
        assert torch.equal(ids, ids2) and torch.equal(emb, emb2) and torch.equal(loss, loss2)


@pytest.mark.parametrize("B,mode", [(65536, O.MODE_ROTATION_TRICK), (200000, O.MODE_ROTATION_TRICK),
                                    (200000, O.MODE_EVAL), (151297, O.MODE_STE)])
def test_large_batches_full_outputs_vs_oracle(B, mode):
    """The bench shape (65,536 rows: one tile per warpgroup) and batches past 148 * 4 * 128 = 75,776 rows, where a
    warpgroup of the d=32 tcgen05 kernel takes a second tile (next-tile prefetch into the residual tile, barrier
    parities of the second use): ids, embeddings, residuals and loss against the oracle on the same rows."""
    D, K, L = 32, 256, 3
    g = torch.Generator().manual_seed(B)
    x = torch.randn(B, D, generator=g)
    cbs = synthetic_codebooks(x[:20000], L, K, seed=4)
    ref = O.rq_forward(x, cbs, mode)
    emb, res, ids, loss = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), mode, beta=0.25)
    assert R.lib().rq_debug_tc_timeout_flag() == 0
    chk = check_ids(ids, ref, cbs, f"d32 3x256 B={B} mode={mode}")
    assert chk.n_mismatch_rows <= 2 + B // 5000        # every one of them inside the tie band (check_ids)
    ok = matching_rows(ids, ref)
    assert_close(emb.cpu()[:, ok], ref.embeddings[:, ok], "emb")
    assert_close(res.cpu()[:, ok], ref.residuals[:, ok], "res")
    assert_close(loss.cpu()[ok], ref.quantize_loss[ok], "loss")
    # and the ids-only launch (tokenisation) of the same batch
    _, _, ids2, _ = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), O.MODE_EVAL, want_embeddings=False,
                                        want_residuals=False, want_loss=False)
    if mode == O.MODE_EVAL:
        assert torch.equal(ids, ids2)


@pytest.mark.parametrize("mode", [O.MODE_ROTATION_TRICK, O.MODE_STE])
@pytest.mark.parametrize("which", ["loss_only", "emb_only", "res_only", "emb_res_loss"])
def test_backward_with_partial_upstream_gradients(which, mode):
    """Outputs nobody differentiated reach the autograd Function as None (no zero tensors are
    materialised); every subset of {embeddings, residuals, loss} must give the oracle's gradients —
    the residuals case goes through the ring-buffered kernel, the others through the default one."""
    B, D, K, L = 3000, 32, 256, 3
    g = torch.Generator().manual_seed(21)
    x = torch.randn(B, D, generator=g)
    cbs = synthetic_codebooks(torch.randn(4 * K, D, generator=g), L, K, seed=3)
    g_emb = torch.randn(L, B, D, generator=g) if "emb" in which else None
    g_res = torch.randn(L, B, D, generator=g) * 0.1 if "res" in which else None
    g_loss = torch.rand(B, generator=g) if "loss" in which else None
    z = torch.zeros(L, B, D)
    ref, gx_ref, gcb_ref = O.rq_forward_backward(x, cbs, mode, 0.25, 0.2, g_emb if g_emb is not None else z,
                                                 g_loss if g_loss is not None else torch.zeros(B), g_res=g_res)
    xin = x.to(DEV).requires_grad_(True)
    cin = torch.stack(cbs).to(DEV).requires_grad_(True)
    emb, res, ids, loss = RF.residual_quantize(xin, cin, mode, 0.25, 0.2)
    outs, grads = [], []
    for o, go in ((emb, g_emb), (res, g_res), (loss, g_loss)):
        if go is not None:
            outs.append(o)
            grads.append(go.to(DEV))
    torch.autograd.backward(outs, grads)
    ok = matching_rows(ids, ref)
    assert ok.float().mean() > 0.995
    assert_close(xin.grad.cpu()[ok], gx_ref[ok], f"{which} grad x")
    if bool(ok.all()):
        assert_close(cin.grad.cpu(), gcb_ref, f"{which} grad codebooks")


@pytest.mark.parametrize("name,kw", [
    ("rqvae_mlp_small.npz", dict(input_dim=64, embed_dim=32, hidden_dims=[48, 40], codebook_size=64, n_layers=3,
                                 n_cat_features=0, mode="rot")),
    ("rqvae_forward_cat18.npz", dict(input_dim=58, embed_dim=16, hidden_dims=[32], codebook_size=32, n_layers=3,
                                     n_cat_features=18, mode="ste")),
])
def test_rqvae_forward_golden(name, kw):
    """`RqVae.forward` (reference modules/rqvae.py:140-165) — loss, reconstruction_loss, rqvae_loss, embs_norm,
    p_unique_ids and the gradients it sends to the encoder and the codebooks — against fixtures produced by the
    un-compiled body of the reference's forward, with n_cat_features = 0 and 18."""
    z = load_golden(name)
    mode = {"rot": R.QuantizeForwardMode.ROTATION_TRICK, "ste": R.QuantizeForwardMode.STE}[kw.pop("mode")]
    m = R.RqVae(codebook_kmeans_init=False, codebook_mode=mode, commitment_weight=0.25, **kw).to(DEV)
    m.load_state_dict({k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")})
    m.train()
    x = torch.from_numpy(z["x"]).to(DEV)
    batch = R.SeqBatch(user_ids=None, ids=None, ids_fut=None, x=x, x_fut=None, seq_mask=None)
    out = m(batch, 0.2)
    assert isinstance(out, R.RqVaeComputedLosses)
    assert_close(out.loss, torch.from_numpy(z["train_loss"]), "loss")
    assert_close(out.reconstruction_loss, torch.from_numpy(z["train_recon"]), "reconstruction_loss")
    assert_close(out.rqvae_loss, torch.from_numpy(z["train_rqvae_loss"]), "rqvae_loss")
    assert out.embs_norm.shape == tuple(z["train_embs_norm"].shape)
    assert_close(out.embs_norm, torch.from_numpy(z["train_embs_norm"]), "embs_norm")
    assert float(out.p_unique_ids) == pytest.approx(float(z["train_p_unique"]), abs=1e-7)
    out.loss.backward()
    assert_close(m.encoder.mlp[0].weight.grad, torch.from_numpy(z["train_grad_enc0"]), "grad encoder.mlp.0")
    gcb = torch.stack([layer.embedding.weight.grad for layer in m.layers])
    assert_close(gcb, torch.from_numpy(z["train_grad_cb"]), "grad codebooks", tol=2e-3)


def test_reference_checkpoint_ids_on_gpu():
    """A checkpoint file written by the reference's driver (train_rqvae.py:209-221) loaded through
    `load_pretrained`: the semantic ids of the probe batch are the reference's."""
    import os
    from conftest import GOLDEN
    z = load_golden("rqvae_checkpoint_ref.npz")
    m = R.RqVae(input_dim=24, embed_dim=8, hidden_dims=[16], codebook_size=16, codebook_kmeans_init=False,
                codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3, n_cat_features=0).to(DEV)
    m.load_pretrained(os.path.join(GOLDEN, "rqvae_checkpoint_ref.pt"))
    m.eval()
    with torch.no_grad():
        out = m.get_semantic_ids(torch.from_numpy(z["x"]).to(DEV))
    same = (out.sem_ids.cpu() == torch.from_numpy(z["eval_ids"])).all(dim=1)
    assert same.float().mean() > 0.97
    assert_close(out.quantize_loss.cpu()[same], torch.from_numpy(z["eval_loss"])[same], "loss")


def test_c1_full_corpus_get_semantic_ids():
    """BASELINE config 1 at full size: 12,101 items, 768-d, encoder [512,256,128] -> d=32, 3x256, codebooks
    k-means-initialised on first use (quantize.py:102-103).  `RqVae.get_semantic_ids` on the GPU against the
    oracle fed with the same latents and codebooks; the torch encoder is checked against its CPU copy."""
    import copy
    N = 12101
    g = torch.Generator().manual_seed(0)
    x = torch.nn.functional.normalize(torch.randn(N, 768, generator=g), dim=1)
    torch.manual_seed(0)
    model = R.RqVae(input_dim=768, embed_dim=32, hidden_dims=[512, 256, 128], codebook_size=256,
                    codebook_kmeans_init=True, codebook_mode=R.QuantizeForwardMode.ROTATION_TRICK, n_layers=3,
                    n_cat_features=0).to(DEV)
    model.eval()
    np.random.seed(0)
    with torch.no_grad():
        out = model.get_semantic_ids(x.to(DEV), 0.001)          # first call: per-level k-means init, then quantise
        assert all(layer.kmeans_initted for layer in model.layers)
        out2 = model.get_semantic_ids(x.to(DEV), 0.001)         # fused path, same codebooks
        lat = model.encode(x.to(DEV))
        cbs = model.effective_codebooks()
    assert out.sem_ids.shape == (N, 3) and out.embeddings.shape == (N, 32, 3)
    assert torch.equal(out.sem_ids, out2.sem_ids)
    ref = O.rq_forward(lat.cpu(), list(cbs.cpu()), O.MODE_EVAL)
    ids_lb = out2.sem_ids.t().contiguous()
    n_mis, n_out1 = check_ids(ids_lb, ref, list(cbs.cpu()), "C1")
    assert n_mis <= 2
    ok = matching_rows(ids_lb, ref)
    assert_close(out2.embeddings.permute(2, 0, 1).cpu()[:, ok], ref.embeddings[:, ok], "emb")
    assert_close(out2.residuals.permute(2, 0, 1).cpu()[:, ok], ref.residuals[:, ok], "residuals")
    assert_close(out2.quantize_loss.cpu()[ok], ref.quantize_loss[ok], "loss")
    lat_cpu = copy.deepcopy(model.encoder).cpu()(x)
    assert_close(lat.cpu(), lat_cpu.detach(), "encoder latents (torch GPU vs torch CPU)", tol=1e-4)
    # k-means-initialised codebooks are all in use at level 0 and most tuples are distinct
    assert int(torch.unique(out2.sem_ids[:, 0]).numel()) >= 250
    assert int(torch.unique(out2.sem_ids, dim=0).shape[0]) > 0.9 * N


def test_full_size_properties_c5():
    """BASELINE config 5 shape (4 levels x 1024 codes, d=64, 65,536 items per GPU), ROTATION_TRICK and
    GUMBEL_SOFTMAX (in-kernel Philox noise): size-independent invariants of forward and backward."""
    B, D, K, L = 65536, 64, 1024, 4
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, D, generator=g)
    cbs = torch.stack(synthetic_codebooks(x[:20000], L, K, seed=6)).to(DEV)
    xd = x.to(DEV)
    g_emb = (torch.randn(L, B, D, generator=g) * 0.01).to(DEV)
    g_loss = torch.full((B,), 1.0 / B, device=DEV)
    _, _, ids_eval, _ = RF.quantize_forward(xd, cbs, O.MODE_EVAL)
    for mode in (O.MODE_ROTATION_TRICK, O.MODE_GUMBEL_SOFTMAX):
        kw = dict(seed=11, offset=3) if mode == O.MODE_GUMBEL_SOFTMAX else {}
        emb, res, ids, loss = RF.quantize_forward(xd, cbs, mode, 0.25, 0.2, **kw)
        assert int(ids.min()) >= 0 and int(ids.max()) < K
        assert torch.equal(res[0], xd)
        assert torch.equal(ids[0], ids_eval[0])              # ids are argmin(dist) in every mode (quantize.py:121)
        for l in range(L - 1):                               # residual chain, rqvae.py:129 — exact
            assert torch.equal(res[l + 1], res[l] - emb[l])
        if mode == O.MODE_ROTATION_TRICK:
            assert torch.equal(ids, ids_eval)
            e = torch.stack([cbs[l][ids[l]] for l in range(L)])
            assert_close(emb, e, "rotation value ~ codeword", tol=1e-4)
        else:
            e = emb                                          # Gumbel: the loss is taken against the soft embedding
            # the soft embedding is a convex combination of the level's codewords
            lo, hi = cbs.min(dim=1).values[:, None, :], cbs.max(dim=1).values[:, None, :]
            assert bool(((emb >= lo - 1e-4) & (emb <= hi + 1e-4)).all())
        want = sum(1.25 * ((res[l] - e[l]) ** 2).sum(-1) for l in range(L))
        assert_close(loss, want, "loss identity", tol=1e-4)
        emb2, _, ids2, loss2 = RF.quantize_forward(xd, cbs, mode, 0.25, 0.2, **kw)
        assert torch.equal(ids, ids2) and torch.equal(emb, emb2) and torch.equal(loss, loss2)
        g_x, g_cb = RF.quantize_backward(g_emb, None, g_loss, res, ids, cbs, mode, 0.25, 0.2, **kw)
        assert bool(torch.isfinite(g_x).all()) and bool(torch.isfinite(g_cb).all())
        assert float(g_cb.abs().sum()) > 0 and float(g_x.abs().sum()) > 0
        if mode == O.MODE_ROTATION_TRICK:
            # codes nobody selected receive no gradient
            for l in range(L):
                unused = torch.ones(K, dtype=torch.bool, device=DEV)
                unused[ids[l]] = False
                assert float(g_cb[l][unused].abs().sum()) == 0.0
