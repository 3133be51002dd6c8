"""GPU: the torch.library registration (rq_vae_recommender_b200/ops.py) and `RqVae.forward` under
`torch.compile(mode="reduce-overhead")`, the decorator the reference puts on it (modules/rqvae.py:140)."""
import pytest
import torch

import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import functional as RF
from rq_vae_recommender_b200 import ops
from conftest import load_golden
from parity_utils import assert_close

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(B=4096, D=32, K=256, L=3, seed=3):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, D, generator=g).to(DEV)
    cb = torch.stack([torch.randn(K, D, generator=g) * (0.6 ** l) for l in range(L)]).to(DEV)
    g_emb = (torch.randn(L, B, D, generator=g) * 0.01).to(DEV)
    g_loss = torch.full((B,), 1.0 / B, device=DEV)
    return x, cb, g_emb, g_loss


@pytest.mark.parametrize("mode", [RF.MODE_ROTATION_TRICK, RF.MODE_STE])
def test_custom_op_equals_autograd_function(mode):
    """torch.ops.rq_engine.quantize_forward (+ its registered autograd formula) against ResidualQuantizeFn: the same
    two launches, so outputs are bit-identical and gradients agree to the order of the codebook atomics."""
    x, cb, g_emb, g_loss = _inputs()
    xa, ca = x.clone().requires_grad_(True), cb.clone().requires_grad_(True)
    emb_a, res_a, ids_a, loss_a = RF.ResidualQuantizeFn.apply(xa, ca, mode, 0.25, 0.2, RF.DIST_L2, None, 0, 0, RF.IMPL_AUTO)
    torch.autograd.backward([emb_a, loss_a], [g_emb, g_loss])
    xb, cbb = x.clone().requires_grad_(True), cb.clone().requires_grad_(True)
    emb_b, res_b, ids_b, loss_b = torch.ops.rq_engine.quantize_forward(xb, cbb, mode, 0.25, 0.2, RF.DIST_L2, None, 0, 0,
                                                                       RF.IMPL_AUTO)
    assert not ids_b.requires_grad and emb_b.requires_grad and loss_b.requires_grad
    torch.autograd.backward([emb_b, loss_b], [g_emb, g_loss])
    assert torch.equal(ids_a, ids_b) and torch.equal(emb_a, emb_b) and torch.equal(res_a, res_b)
    assert torch.equal(loss_a, loss_b)
    assert torch.equal(xa.grad, xb.grad)
    assert_close(cbb.grad, ca.grad, "g_codebooks", tol=1e-5)


def test_opcheck_schema_and_fake():
    x, cb, _, _ = _inputs(B=1024)
    args = (x, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2, RF.DIST_L2, None, 0, 0, RF.IMPL_AUTO)
    torch.library.opcheck(torch.ops.rq_engine.quantize_forward.default, args,
                          test_utils=("test_schema", "test_faketensor"))
    _, res, ids, _ = RF.quantize_forward(x, cb, RF.MODE_ROTATION_TRICK)
    torch.library.opcheck(torch.ops.rq_engine.corpus_table.default, (ids.t(), 256),
                          test_utils=("test_schema", "test_faketensor"))
    assert set(ops.OPS) <= set(dir(torch.ops.rq_engine))


def _forward_fixture(name, kw):
    z = load_golden(name)
    kw = dict(kw)
    mode = {"rot": R.QuantizeForwardMode.ROTATION_TRICK, "ste": R.QuantizeForwardMode.STE}[kw.pop("mode")]
    m = R.RqVae(codebook_kmeans_init=False, codebook_mode=mode, commitment_weight=0.25, **kw).to(DEV)
    m.load_state_dict({k[4:]: torch.from_numpy(z[k]) for k in z.keys() if k.startswith("sd::")})
    m.train()
    x = torch.from_numpy(z["x"]).to(DEV)
    return z, m, R.SeqBatch(user_ids=None, ids=None, ids_fut=None, x=x, x_fut=None, seq_mask=None)


@pytest.mark.parametrize("name,kw", [
    ("rqvae_mlp_small.npz", dict(input_dim=64, embed_dim=32, hidden_dims=[48, 40], codebook_size=64, n_layers=3,
                                 n_cat_features=0, mode="rot")),
    ("rqvae_forward_cat18.npz", dict(input_dim=58, embed_dim=16, hidden_dims=[32], codebook_size=32, n_layers=3,
                                     n_cat_features=18, mode="ste")),
])
def test_rqvae_forward_compiled_golden(name, kw, monkeypatch):
    """`RqVae.forward` with the reference's decorator active: the compiled step (Dynamo through the custom ops,
    AOTAutograd backward, CUDA-graph replay) reproduces the reference's fixture, on the first and on a replayed call."""
    monkeypatch.setattr(R.RqVae, "compile_forward", True)
    z, m, batch = _forward_fixture(name, kw)
    for rep in range(3):                                # compile, CUDA-graph recording, replay
        m.zero_grad(set_to_none=True)
        with _no_compile_warning():
            out = m(batch, 0.2)
        loss = out.loss.clone()
        assert_close(loss, torch.from_numpy(z["train_loss"]), "loss")
        assert_close(out.reconstruction_loss.clone(), torch.from_numpy(z["train_recon"]), "reconstruction_loss")
        assert_close(out.rqvae_loss.clone(), torch.from_numpy(z["train_rqvae_loss"]), "rqvae_loss")
        assert_close(out.embs_norm.clone(), torch.from_numpy(z["train_embs_norm"]), "embs_norm")
        assert float(out.p_unique_ids) == pytest.approx(float(z["train_p_unique"]), abs=1e-7)
        out.loss.backward()
        assert_close(m.encoder.mlp[0].weight.grad.clone(), torch.from_numpy(z["train_grad_enc0"]), "grad encoder.mlp.0")
        gcb = torch.stack([layer.embedding.weight.grad.clone() for layer in m.layers])
        assert_close(gcb, torch.from_numpy(z["train_grad_cb"]), "grad codebooks", tol=2e-3)
    from rq_vae_recommender_b200.modules.rqvae import _COMPILED
    assert _COMPILED[m].get("ok") and not _COMPILED[m].get("failed")     # it really ran compiled, not the eager fallback


class _no_compile_warning:
    """Fails the test when RqVae.forward fell back to eager (it warns when torch.compile raised)."""

    def __enter__(self):
        import warnings
        self._cm = warnings.catch_warnings(record=True)
        self._w = self._cm.__enter__()
        warnings.simplefilter("always")
        return self

    def __exit__(self, *a):
        self._cm.__exit__(*a)
        bad = [str(w.message) for w in self._w if "torch.compile failed" in str(w.message)]
        assert not bad, bad
        return False
