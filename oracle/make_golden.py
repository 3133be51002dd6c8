"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

    python oracle/make_golden.py            # needs /root/reference (read-only mount)

The reference ships no golden vectors, so the oracle is pinned against
outputs of the reference itself: this script imports modules.quantize.Quantize,
modules.rqvae.RqVae, init.kmeans.Kmeans and modules.tokenizer.semids.SemanticIdTokenizer
from /root/reference (behind the name-only stubs in oracle/_stubs), drives them on seeded
inputs on CPU (true fp32) and stores inputs + outputs.  The fixtures travel to the GPU box;
the reference does not.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_import  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def make_codebooks(L, K, D, gen, scale0=1.0):
    return [torch.randn(K, D, generator=gen) * (scale0 * 0.55 ** l) for l in range(L)]


def golden_quantize_level(ref, name, B, D, K, seed):
    """One Quantize layer (modules/quantize.py:99-156), all four branches, with autograd."""
    Q = ref.quantize
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, D, generator=gen)
    cb = torch.randn(K, D, generator=gen)
    # make a few exact duplicates so that "first index wins" (quantize.py:121) is exercised
    cb[7] = cb[3]
    x[5] = cb[3] + 1e-3
    g_emb = torch.randn(B, D, generator=gen)
    g_loss = torch.rand(B, generator=gen)
    out = {"x": _np(x), "codebook": _np(cb), "g_emb": _np(g_emb), "g_loss": _np(g_loss),
           "beta": np.float32(0.25), "temperature": np.float32(0.2)}
    for mode_name, mode, training in [("eval", Q.QuantizeForwardMode.STE, False),
                                      ("ste", Q.QuantizeForwardMode.STE, True),
                                      ("rot", Q.QuantizeForwardMode.ROTATION_TRICK, True),
                                      ("gumbel", Q.QuantizeForwardMode.GUMBEL_SOFTMAX, True)]:
        layer = Q.Quantize(embed_dim=D, n_embed=K, do_kmeans_init=False, forward_mode=mode,
                           commitment_weight=0.25)
        with torch.no_grad():
            layer.embedding.weight.copy_(cb)
        layer.train(training)
        xin = x.clone().requires_grad_(True)
        torch.manual_seed(seed + 17)
        res = layer(xin, temperature=0.2)
        if mode_name == "gumbel":
            # the uniform draw the layer consumed: torch.rand(dist.shape) right after the seed
            torch.manual_seed(seed + 17)
            out["gumbel_uniform"] = _np(torch.rand(B, K))
        obj = (res.embeddings * g_emb).sum() + (res.loss * g_loss).sum()
        obj.backward()
        out[f"{mode_name}_emb"] = _np(res.embeddings)
        out[f"{mode_name}_ids"] = _np(res.ids)
        out[f"{mode_name}_loss"] = _np(res.loss)
        out[f"{mode_name}_gx"] = _np(xin.grad)
        gw = layer.embedding.weight.grad
        out[f"{mode_name}_gcb"] = _np(gw if gw is not None else torch.zeros_like(cb))
    # COSINE distance (quantize.py:114-117), eval branch
    layer = Q.Quantize(embed_dim=D, n_embed=K, do_kmeans_init=False,
                       forward_mode=Q.QuantizeForwardMode.STE,
                       distance_mode=Q.QuantizeDistance.COSINE)
    with torch.no_grad():
        layer.embedding.weight.copy_(cb)
    layer.eval()
    res = layer(x, temperature=0.2)
    out["cosine_ids"] = _np(res.ids)
    out["cosine_emb"] = _np(res.embeddings)
    out["cosine_loss"] = _np(res.loss)
    np.savez_compressed(os.path.join(OUT, name), **out)


def golden_rq_stack(ref, name, B, D, K, L, seed, modes=("eval", "ste", "rot", "gumbel"),
                    normalize=False, sim_vq=False):
    """RqVae.get_semantic_ids (modules/rqvae.py:114-138) with an identity encoder
    (hidden_dims=[], weight = I) so that the level loop runs on given latents, plus autograd
    on embeddings [B,D,L] and quantize_loss."""
    Q, R = ref.quantize, ref.rqvae
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, D, generator=gen)
    if normalize:
        x = torch.nn.functional.normalize(x, dim=-1)
    cbs = make_codebooks(L, K, D, gen)
    g_emb = torch.randn(L, B, D, generator=gen)
    g_loss = torch.rand(B, generator=gen)
    out = {"x": _np(x), "codebooks": _np(torch.stack(cbs)), "g_emb": _np(g_emb), "g_loss": _np(g_loss),
           "beta": np.float32(0.25), "temperature": np.float32(0.2),
           "normalize": np.bool_(normalize), "sim_vq": np.bool_(sim_vq)}
    mode_map = {"eval": (Q.QuantizeForwardMode.STE, False), "ste": (Q.QuantizeForwardMode.STE, True),
                "rot": (Q.QuantizeForwardMode.ROTATION_TRICK, True),
                "gumbel": (Q.QuantizeForwardMode.GUMBEL_SOFTMAX, True)}
    sim_w = None
    for mode_name in modes:
        mode, training = mode_map[mode_name]
        torch.manual_seed(seed + 1)
        model = R.RqVae(input_dim=D, embed_dim=D, hidden_dims=[], codebook_size=K,
                        codebook_kmeans_init=False, codebook_normalize=normalize,
                        codebook_sim_vq=sim_vq, codebook_mode=mode, n_layers=L,
                        commitment_weight=0.25, n_cat_features=0)
        with torch.no_grad():
            model.encoder.mlp[0].weight.copy_(torch.eye(D))
            for l in range(L):
                model.layers[l].embedding.weight.copy_(cbs[l])
            if sim_vq:
                sim_w = torch.stack([model.layers[l].out_proj[0].weight.detach().clone() for l in range(L)])
        model.train(training)
        xin = x.clone().requires_grad_(True)
        torch.manual_seed(seed + 29)
        res = model.get_semantic_ids(xin, gumbel_t=0.2)
        if mode_name == "gumbel":
            torch.manual_seed(seed + 29)
            out["gumbel_uniform"] = _np(torch.stack([torch.rand(B, K) for _ in range(L)]))
        # reference layouts: embeddings/residuals are [B,D,L] views of [L,B,D] memory
        emb_lbd = res.embeddings.permute(2, 0, 1)
        res_lbd = res.residuals.permute(2, 0, 1)
        obj = (emb_lbd * g_emb).sum() + (res.quantize_loss * g_loss).sum()
        obj.backward()
        out[f"{mode_name}_emb"] = _np(emb_lbd)
        out[f"{mode_name}_res"] = _np(res_lbd)
        out[f"{mode_name}_ids"] = _np(res.sem_ids.T)
        out[f"{mode_name}_loss"] = _np(res.quantize_loss)
        out[f"{mode_name}_gx"] = _np(xin.grad)
        out[f"{mode_name}_gcb"] = _np(torch.stack([
            (model.layers[l].embedding.weight.grad
             if model.layers[l].embedding.weight.grad is not None else torch.zeros(K, D))
            for l in range(L)]))
        out[f"{mode_name}_emb_strides"] = np.asarray(res.embeddings.stride(), dtype=np.int64)
        out[f"{mode_name}_ids_strides"] = np.asarray(res.sem_ids.stride(), dtype=np.int64)
    if sim_w is not None:
        out["sim_vq_weights"] = _np(sim_w)
    np.savez_compressed(os.path.join(OUT, name), **out)


def golden_rqvae_mlp(ref, name, seed):
    """RqVae with a real (small) encoder MLP: end-to-end x -> semantic ids, eval branch, and
    the full training forward losses (modules/rqvae.py:141-165, called un-compiled)."""
    Q, R = ref.quantize, ref.rqvae
    torch.manual_seed(seed)
    B, IN, D, K, L = 300, 64, 32, 64, 3
    model = R.RqVae(input_dim=IN, embed_dim=D, hidden_dims=[48, 40], codebook_size=K,
                    codebook_kmeans_init=False, codebook_normalize=False, codebook_sim_vq=False,
                    codebook_mode=Q.QuantizeForwardMode.ROTATION_TRICK, n_layers=L,
                    commitment_weight=0.25, n_cat_features=0)
    gen = torch.Generator().manual_seed(seed + 1)
    x = torch.nn.functional.normalize(torch.randn(B, IN, generator=gen), dim=-1)
    with torch.no_grad():
        lat = model.encode(x)
        res = lat.clone()
        for l in range(L):
            idx = torch.randperm(B, generator=gen)[:K]
            model.layers[l].embedding.weight.copy_(res[idx] * 0.9)
            model.eval()
            res = res - model.layers[l](res, 0.2).embeddings
    model.eval()
    with torch.no_grad():
        ev = model.get_semantic_ids(x)
    out = {"x": _np(x)}
    for k, v in model.state_dict().items():
        out["sd::" + k] = _np(v)
    out["eval_ids"] = _np(ev.sem_ids)
    out["eval_loss"] = _np(ev.quantize_loss)
    out["eval_emb"] = _np(ev.embeddings.permute(2, 0, 1))
    # training forward, un-compiled body of RqVae.forward
    model.train()
    fwd = R.RqVae.forward
    inner = getattr(fwd, "_torchdynamo_orig_callable", None) or getattr(fwd, "__wrapped__", None)
    batch = ref.schemas.SeqBatch(user_ids=None, ids=None, ids_fut=None, x=x, x_fut=None, seq_mask=None)
    if inner is not None:
        losses = inner(model, batch, 0.2)
        out["train_loss"] = _np(losses.loss)
        out["train_recon"] = _np(losses.reconstruction_loss)
        out["train_rqvae_loss"] = _np(losses.rqvae_loss)
        out["train_p_unique"] = _np(losses.p_unique_ids)
        out["train_embs_norm"] = _np(losses.embs_norm)
        losses.loss.backward()
        out["train_grad_enc0"] = _np(model.encoder.mlp[0].weight.grad)
        out["train_grad_cb"] = _np(torch.stack([model.layers[l].embedding.weight.grad for l in range(L)]))
    np.savez_compressed(os.path.join(OUT, name), **out)


def golden_rqvae_forward_cat(ref, name, seed, n_cat=18):
    """`RqVae.forward` (modules/rqvae.py:140-165, un-compiled body) with categorical features: the decoder output is
    split into an L2-normalised dense part and n_cat logits (rqvae.py:146), CategoricalReconstuctionLoss
    (modules/loss.py:19-30).  All five outputs + gradients."""
    Q, R = ref.quantize, ref.rqvae
    torch.manual_seed(seed)
    B, IN, D, K, L = 160, 40 + n_cat, 16, 32, 3
    model = R.RqVae(input_dim=IN, embed_dim=D, hidden_dims=[32], codebook_size=K, codebook_kmeans_init=False,
                    codebook_mode=Q.QuantizeForwardMode.STE, n_layers=L, commitment_weight=0.25,
                    n_cat_features=n_cat)
    gen = torch.Generator().manual_seed(seed + 1)
    dense = torch.nn.functional.normalize(torch.randn(B, IN - n_cat, generator=gen), dim=-1)
    cats = (torch.rand(B, n_cat, generator=gen) < 0.2).float()
    x = torch.cat([dense, cats], dim=-1)
    with torch.no_grad():
        res = model.encode(x).clone()
        model.eval()
        for l in range(L):
            idx = torch.randint(0, B, (K,), generator=gen)         # duplicates on purpose: p_unique_ids < 1
            model.layers[l].embedding.weight.copy_(res[idx] * 0.9)
            res = res - model.layers[l](res, 0.2).embeddings
    x[B // 2:B // 2 + 20] = x[:20]                                   # repeated items -> repeated id tuples
    out = {"x": _np(x), "n_cat": np.int64(n_cat)}
    for k, v in model.state_dict().items():
        out["sd::" + k] = _np(v)
    model.train()
    fwd = R.RqVae.forward
    inner = getattr(fwd, "_torchdynamo_orig_callable", None) or getattr(fwd, "__wrapped__", None)
    assert inner is not None, "cannot reach the un-compiled body of RqVae.forward"
    batch = ref.schemas.SeqBatch(user_ids=None, ids=None, ids_fut=None, x=x, x_fut=None, seq_mask=None)
    losses = inner(model, batch, 0.2)
    out["train_loss"] = _np(losses.loss)
    out["train_recon"] = _np(losses.reconstruction_loss)
    out["train_rqvae_loss"] = _np(losses.rqvae_loss)
    out["train_p_unique"] = _np(losses.p_unique_ids)
    out["train_embs_norm"] = _np(losses.embs_norm)
    losses.loss.backward()
    out["train_grad_enc0"] = _np(model.encoder.mlp[0].weight.grad)
    out["train_grad_cb"] = _np(torch.stack([model.layers[l].embedding.weight.grad for l in range(L)]))
    np.savez_compressed(os.path.join(OUT, name), **out)


def golden_checkpoint(ref, name, seed):
    """A checkpoint file written exactly as the reference's training driver writes it (train_rqvae.py:209-221:
    {"iter", "model", "model_config", "optimizer"}; model_config = RqVae.config = locals() of __init__, i.e. it
    pickles the reference's own classes), plus the ids the reference computes from it on a probe batch."""
    Q, R = ref.quantize, ref.rqvae
    torch.manual_seed(seed)
    B, IN, D, K, L = 64, 24, 8, 16, 3
    model = R.RqVae(input_dim=IN, embed_dim=D, hidden_dims=[16], codebook_size=K, codebook_kmeans_init=False,
                    codebook_mode=Q.QuantizeForwardMode.ROTATION_TRICK, n_layers=L, commitment_weight=0.25,
                    n_cat_features=0)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
    gen = torch.Generator().manual_seed(seed + 1)
    x = torch.nn.functional.normalize(torch.randn(B, IN, generator=gen), dim=-1)
    model.train()
    out_t = model.get_semantic_ids(x, 0.2)
    (out_t.quantize_loss.mean() + out_t.embeddings.sum() * 1e-3).backward()
    opt.step()                                                        # so that the optimizer state is not empty
    state = {"iter": 41, "model": model.state_dict(), "model_config": model.config, "optimizer": opt.state_dict()}
    torch.save(state, os.path.join(OUT, name + ".pt"))
    model.eval()
    with torch.no_grad():
        ev = model.get_semantic_ids(x)
    out = {"x": _np(x), "eval_ids": _np(ev.sem_ids), "eval_loss": _np(ev.quantize_loss), "iter": np.int64(41)}
    for k, v in model.state_dict().items():
        out["sd::" + k] = _np(v)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)


def golden_kmeans(ref, name, seed):
    """init/kmeans.py: a full run (seeded np.random.choice init, :36) and single
    _update_centroids steps, one of them with empty clusters (torch.randint reseed, :53)."""
    KM = ref.kmeans
    gen = torch.Generator().manual_seed(seed)
    B, D, K = 600, 8, 16
    x = torch.randn(B, D, generator=gen) + 3.0 * torch.randn(6, D, generator=gen)[torch.randint(0, 6, (B,), generator=gen)]
    np.random.seed(seed)
    init_idx = np.random.choice(B, K, replace=False)
    np.random.seed(seed)
    km = KM.Kmeans(k=K)
    res = km.run(x)
    out = {"x": _np(x), "init_idx": init_idx.astype(np.int64),
           "run_centroids": _np(res.centroids), "run_assignment": _np(res.assignment)}
    # single step from given centroids, no empties
    km2 = KM.Kmeans(k=K)
    c0 = x[torch.as_tensor(init_idx)].clone()
    km2.centroids = c0.clone()
    km2._update_centroids(x)
    out["step_c0"] = _np(c0)
    out["step_c1"] = _np(km2.centroids)
    out["step_assign"] = _np(km2.assignment)
    # single step with two far-away (hence empty) centroids
    c_e = c0.clone()
    c_e[2] = 1e3
    c_e[9] = -1e3
    km3 = KM.Kmeans(k=K)
    km3.centroids = c_e.clone()
    torch.manual_seed(seed + 5)
    km3._update_centroids(x)
    torch.manual_seed(seed + 5)
    rows = [int(torch.randint(0, B, (1,))) for _ in range(2)]
    out["empty_c0"] = _np(c_e)
    out["empty_c1"] = _np(km3.centroids)
    out["empty_assign"] = _np(km3.assignment)
    out["empty_reseed_rows"] = np.asarray(rows, dtype=np.int64)
    # kmeans_init_ (kmeans.py:8-15)
    w = torch.zeros(K, D)
    np.random.seed(seed)
    KM.kmeans_init_(w, x)
    out["init_tensor"] = _np(w)
    np.savez_compressed(os.path.join(OUT, name), **out)


class _FakeItemData(torch.utils.data.Dataset):
    """Mimics data/processed.py:74-84 (`ItemData.__getitem__` on a list of indices)."""

    def __init__(self, x, SeqBatch):
        self.item_data = x
        self.SeqBatch = SeqBatch

    def __len__(self):
        return self.item_data.shape[0]

    def __getitem__(self, idx):
        item_ids = torch.tensor(idx).unsqueeze(0) if not isinstance(idx, torch.Tensor) else idx
        x = self.item_data[idx, :768]
        return self.SeqBatch(
            user_ids=-1 * torch.ones_like(item_ids.squeeze(0)), ids=item_ids,
            ids_fut=-1 * torch.ones_like(item_ids.squeeze(0)), x=x,
            x_fut=-1 * torch.ones_like(item_ids.squeeze(0)),
            seq_mask=torch.ones_like(item_ids, dtype=bool))


def golden_tokenizer(ref, name, seed):
    """modules/tokenizer/semids.py: precompute_corpus_ids (3 batches of 512, many duplicate
    tuples), exists_prefix (with its floor-division quirk, :114) and the cached-gather forward."""
    S = ref.semids
    torch.manual_seed(seed)
    N, IN, D, K, L = 1300, 24, 8, 6, 3
    tok = S.SemanticIdTokenizer(input_dim=IN, output_dim=D, hidden_dims=[16], codebook_size=K,
                                n_layers=L, n_cat_feats=0)
    gen = torch.Generator().manual_seed(seed + 1)
    x = torch.randn(N, IN, generator=gen)
    x[700:720] = x[10:30]          # exact duplicate items across batches
    x[600:605] = x[598]            # in-batch duplicates straddling nothing special
    with torch.no_grad():
        for l in range(L):
            tok.rq_vae.layers[l].embedding.weight.copy_(torch.randn(K, D, generator=gen) * (0.5 ** l))
    ds = _FakeItemData(x, ref.schemas.SeqBatch)
    table = tok.precompute_corpus_ids(ds)
    out = {"x": _np(x), "cached_ids": _np(table)}
    for k, v in tok.rq_vae.state_dict().items():
        out["sd::" + k] = _np(v)
    # exists_prefix: [n, m, p]
    pg = torch.Generator().manual_seed(seed + 2)
    for p in (1, 2, 3, 4):
        pref = torch.randint(0, K, (37, 5, p), generator=pg)
        pref[0, 0] = table[3, :p]
        pref[36, 4] = table[100, :p]
        pref[20, 2] = table[900, :p]
        out[f"prefix_q{p}"] = _np(pref)
        out[f"prefix_r{p}"] = _np(tok.exists_prefix(pref))
    # forward from cache (semids.py:137-154)
    B, Ns = 6, 5
    ids = torch.randint(0, N, (B, Ns), generator=pg)
    ids_fut = torch.randint(0, N, (B, 1), generator=pg)
    seq_mask = torch.rand(B, Ns, generator=pg) > 0.3
    batch = ref.schemas.SeqBatch(user_ids=torch.arange(B), ids=ids, ids_fut=ids_fut,
                                 x=torch.zeros(B, Ns, IN), x_fut=torch.zeros(B, 1, IN), seq_mask=seq_mask)
    tk = tok(batch)
    out.update({"fw_ids": _np(ids), "fw_ids_fut": _np(ids_fut), "fw_seq_mask": _np(seq_mask),
                "fw_sem_ids": _np(tk.sem_ids), "fw_sem_ids_fut": _np(tk.sem_ids_fut),
                "fw_seq_mask_out": _np(tk.seq_mask), "fw_token_type_ids": _np(tk.token_type_ids),
                "fw_token_type_ids_fut": _np(tk.token_type_ids_fut)})
    np.savez_compressed(os.path.join(OUT, name), **out)


def main():
    ref = ref_import.load()
    torch.set_num_threads(1)       # fixed reduction order for the committed fixtures
    os.makedirs(OUT, exist_ok=True)
    golden_quantize_level(ref, "quantize_level_d32_k256.npz", B=193, D=32, K=256, seed=1)
    golden_quantize_level(ref, "quantize_level_d64_k1024.npz", B=70, D=64, K=1024, seed=2)
    golden_rq_stack(ref, "rq_stack_d32_k256_l3.npz", B=321, D=32, K=256, L=3, seed=3)
    golden_rq_stack(ref, "rq_stack_d64_k256_l3.npz", B=130, D=64, K=256, L=3, seed=4, modes=("eval", "rot"))
    golden_rq_stack(ref, "rq_stack_d64_k1024_l4.npz", B=48, D=64, K=1024, L=4, seed=5, modes=("rot", "gumbel"))
    golden_rq_stack(ref, "rq_stack_norm_simvq_d16_k32_l2.npz", B=90, D=16, K=32, L=2, seed=6,
                    modes=("eval", "ste", "rot"), normalize=True, sim_vq=True)
    golden_rqvae_mlp(ref, "rqvae_mlp_small.npz", seed=7)
    golden_rqvae_forward_cat(ref, "rqvae_forward_cat18.npz", seed=10)
    golden_checkpoint(ref, "rqvae_checkpoint_ref", seed=11)
    golden_kmeans(ref, "kmeans_small.npz", seed=8)
    golden_tokenizer(ref, "tokenizer_small.npz", seed=9)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
