"""Pins oracle/rq_oracle.py against fixtures generated from the UNMODIFIED reference
(oracle/make_golden.py).  CPU only.  Floats must agree to the last bit where the op
sequence is the same ATen sequence (single-thread fixtures; a multi-thread MKL run may
re-associate, hence the 2e-6 relative fallback), integers bit-exact."""
import numpy as np
import pytest
import torch

import rq_oracle as O
from conftest import load_golden

MODES = {"eval": O.MODE_EVAL, "ste": O.MODE_STE, "rot": O.MODE_ROTATION_TRICK, "gumbel": O.MODE_GUMBEL_SOFTMAX}


def close(a, b, rtol=2e-6, atol=2e-6):
    a = np.asarray(a); b = np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=atol)


@pytest.mark.parametrize("name", ["quantize_level_d32_k256.npz", "quantize_level_d64_k1024.npz"])
def test_quantize_level_all_modes(name):
    torch.set_num_threads(1)
    z = load_golden(name)
    x, cb = torch.from_numpy(z["x"]), torch.from_numpy(z["codebook"])
    g_emb, g_loss = torch.from_numpy(z["g_emb"]), torch.from_numpy(z["g_loss"])
    for mname, mode in MODES.items():
        xin = x.clone().requires_grad_(True)
        cin = cb.clone().requires_grad_(True)
        u = torch.from_numpy(z["gumbel_uniform"]) if mname == "gumbel" else None
        out = O.quantize_level(xin, cin, mode, 0.25, 0.2, O.DIST_L2, u)
        assert np.array_equal(out.ids.numpy(), z[f"{mname}_ids"]), mname
        close(out.emb_out.detach(), z[f"{mname}_emb"])
        close(out.loss.detach(), z[f"{mname}_loss"])
        ((out.emb_out * g_emb).sum() + (out.loss * g_loss).sum()).backward()
        close(xin.grad, z[f"{mname}_gx"], rtol=1e-5, atol=1e-5)
        gc = cin.grad if cin.grad is not None else torch.zeros_like(cb)
        close(gc, z[f"{mname}_gcb"], rtol=1e-5, atol=1e-5)
    # duplicate codeword rows 3 and 7: first index wins (quantize.py:121)
    assert z["eval_ids"][5] == 3
    out = O.quantize_level(x, cb, O.MODE_EVAL, 0.25, 0.2, O.DIST_COSINE)
    assert np.array_equal(out.ids.numpy(), z["cosine_ids"])
    close(out.emb_out, z["cosine_emb"]); close(out.loss, z["cosine_loss"])


@pytest.mark.parametrize("name", ["rq_stack_d32_k256_l3.npz", "rq_stack_d64_k256_l3.npz",
                                  "rq_stack_d64_k1024_l4.npz", "rq_stack_norm_simvq_d16_k32_l2.npz"])
def test_rq_stack(name):
    torch.set_num_threads(1)
    z = load_golden(name)
    x = torch.from_numpy(z["x"])
    w = torch.from_numpy(z["codebooks"])
    L = w.shape[0]
    sim = torch.from_numpy(z["sim_vq_weights"]) if "sim_vq_weights" in z else None
    norm = bool(z["normalize"])
    g_emb, g_loss = torch.from_numpy(z["g_emb"]), torch.from_numpy(z["g_loss"])
    for mname, mode in MODES.items():
        if f"{mname}_ids" not in z:
            continue
        ws = [w[l].clone().requires_grad_(True) for l in range(L)]
        # codebook_normalize applies to level 0 only (modules/rqvae.py:70)
        cbs = [O.effective_codebook(ws[l], None if sim is None else sim[l], norm and l == 0) for l in range(L)]
        xin = x.clone().requires_grad_(True)
        u = torch.from_numpy(z["gumbel_uniform"]) if mname == "gumbel" else None
        # the fixture's encoder is Linear(D, D) with weight I (+ l2norm when codebook_normalize,
        # modules/rqvae.py:76-81), so gradients pass through that normalisation
        lat = O.mlp_forward(xin, [torch.eye(x.shape[1])], normalize=norm)
        out = O.rq_forward(lat, cbs, mode, 0.25, 0.2, O.DIST_L2, u)
        assert np.array_equal(out.sem_ids.numpy(), z[f"{mname}_ids"]), mname
        close(out.embeddings.detach(), z[f"{mname}_emb"])
        close(out.residuals.detach(), z[f"{mname}_res"])
        close(out.quantize_loss.detach(), z[f"{mname}_loss"], rtol=1e-5)
        ((out.embeddings * g_emb).sum() + (out.quantize_loss * g_loss).sum()).backward()
        close(xin.grad, z[f"{mname}_gx"], rtol=2e-5, atol=2e-5)
        gc = torch.stack([c.grad if c.grad is not None else torch.zeros_like(c) for c in ws])
        close(gc, z[f"{mname}_gcb"], rtol=2e-5, atol=2e-5)
        # the reference hands out [B,D,L] / [B,L] views over [L,B,D] / [L,B] memory
        B, D = x.shape
        assert tuple(z[f"{mname}_emb_strides"]) == (D, 1, B * D)
        assert tuple(z[f"{mname}_ids_strides"]) == (1, B)


def test_rqvae_mlp_end_to_end():
    torch.set_num_threads(1)
    z = load_golden("rqvae_mlp_small.npz")
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")}
    x = torch.from_numpy(z["x"])
    enc = [sd[f"encoder.mlp.{i}.weight"] for i in (0, 2, 4)]
    lat = O.mlp_forward(x, enc, normalize=False)
    cbs = [sd[f"layers.{l}.embedding.weight"] for l in range(3)]
    out = O.rq_forward(lat, cbs, O.MODE_EVAL)
    assert np.array_equal(out.sem_ids.T.numpy(), z["eval_ids"])
    close(out.quantize_loss, z["eval_loss"], rtol=1e-5)
    close(out.embeddings, z["eval_emb"])


def test_kmeans():
    torch.set_num_threads(1)
    z = load_golden("kmeans_small.npz")
    x = torch.from_numpy(z["x"])
    c1, a1 = O.kmeans_update(x, torch.from_numpy(z["step_c0"]))
    assert np.array_equal(a1.numpy(), z["step_assign"])
    close(c1, z["step_c1"], rtol=1e-6, atol=1e-6)
    c2, a2 = O.kmeans_update(x, torch.from_numpy(z["empty_c0"]), reseed_rows=z["empty_reseed_rows"].tolist())
    assert np.array_equal(a2.numpy(), z["empty_assign"])
    close(c2, z["empty_c1"], rtol=1e-6, atol=1e-6)
    c, a, _ = O.kmeans_run(x, 16, init_idx=z["init_idx"])
    assert np.array_equal(a.numpy(), z["run_assignment"])
    close(c, z["run_centroids"], rtol=1e-6, atol=1e-6)
    close(c, z["init_tensor"], rtol=1e-6, atol=1e-6)
    with pytest.raises(ValueError):
        O.kmeans_update(torch.zeros(0, 8), torch.zeros(4, 8))


def test_tokenizer_integer_paths():
    z = load_golden("tokenizer_small.npz")
    table = z["cached_ids"]
    assert table.dtype == np.int64
    assert np.array_equal(O.corpus_ids_table(table[:, :-1]), table)
    for p in (1, 2, 3, 4):
        got = O.exists_prefix(table, z[f"prefix_q{p}"], strict_reference_quirk=True)
        assert np.array_equal(got, z[f"prefix_r{p}"]), p
        full = O.exists_prefix(table, z[f"prefix_q{p}"], strict_reference_quirk=False)
        assert np.array_equal(full[:32], z[f"prefix_r{p}"][:32])
        assert full[36, 4]          # planted hit beyond the last full chunk of 16
    sem = O.tokenize_from_cache(table, z["fw_ids"], z["fw_seq_mask"])
    assert np.array_equal(sem, z["fw_sem_ids"])
    assert np.array_equal(O.tokenize_from_cache(table, z["fw_ids_fut"]), z["fw_sem_ids_fut"])
    with pytest.raises(Exception):
        O.exists_prefix(None, z["prefix_q1"])


def test_tokenizer_semantic_ids_from_state_dict():
    """End to end: encoder + 3 levels from the fixture's state dict reproduce the id table."""
    torch.set_num_threads(1)
    z = load_golden("tokenizer_small.npz")
    sd = {k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")}
    lat = O.mlp_forward(torch.from_numpy(z["x"]), [sd["encoder.mlp.0.weight"], sd["encoder.mlp.2.weight"]], False)
    out = O.rq_forward(lat, [sd[f"layers.{l}.embedding.weight"] for l in range(3)], O.MODE_EVAL)
    assert np.array_equal(out.sem_ids.T.numpy(), z["cached_ids"][:, :3])


def test_id_diversity_oracle_known_answer():
    """oracle.id_diversity (train_rqvae.py:227-239) on a table whose answer is known in closed form: tuples with
    multiplicities 3, 2, 1 out of 6 -> entropy = -(1/2 ln 1/2 + 1/3 ln 1/3 + 1/6 ln 1/6)."""
    import math
    ids = np.array([[0, 1], [0, 1], [2, 1], [0, 1], [2, 1], [3, 0]], dtype=np.int64)
    table = torch.from_numpy(O.corpus_ids_table(ids).reshape(6, 3))
    log = O.id_diversity(table, 2, 4)
    want = -(0.5 * math.log(0.5) + (1 / 3) * math.log(1 / 3) + (1 / 6) * math.log(1 / 6))
    assert abs(log["rqvae_entropy"] - want) < 1e-6
    assert log["codebook_usage_0"] == 3 / 4 and log["codebook_usage_1"] == 2 / 4
    assert abs(log["max_id_duplicates"] - 2 / 6) < 1e-7

