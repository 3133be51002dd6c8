"""GPU: Gumbel-softmax backward (rq_quantize_backward, mode 1) against the reference's autograd
(golden fixtures) and the CPU oracle on seeded shapes; Philox-noise mode against finite differences."""
import pytest
import torch

import rq_oracle as O
from rq_vae_recommender_b200 import functional as RF
from conftest import load_golden
from parity_utils import assert_close, synthetic_codebooks

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("name", ["rq_stack_d32_k256_l3.npz", "rq_stack_d64_k256_l3.npz", "rq_stack_d64_k1024_l4.npz"])
def test_golden_gumbel_backward(name):
    """Gradients the unmodified reference produced (autograd through rand->log->softmax->GEMM)."""
    z = load_golden(name)
    if "gumbel_gx" not in z:
        pytest.skip("fixture has no Gumbel gradients")
    x = torch.from_numpy(z["x"]).to(DEV).requires_grad_(True)
    cbs = torch.from_numpy(z["codebooks"]).to(DEV).requires_grad_(True)
    gu = torch.from_numpy(z["gumbel_uniform"]).to(DEV)
    g_emb, g_loss = torch.from_numpy(z["g_emb"]).to(DEV), torch.from_numpy(z["g_loss"]).to(DEV)
    emb, res, ids, loss = RF.residual_quantize(x, cbs, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, RF.DIST_L2, gu)
    ((emb * g_emb).sum() + (loss * g_loss).sum()).backward()
    assert_close(x.grad.cpu(), torch.from_numpy(z["gumbel_gx"]), "gumbel grad x")
    assert_close(cbs.grad.cpu(), torch.from_numpy(z["gumbel_gcb"]), "gumbel grad codebooks")


@pytest.mark.parametrize("B,D,K,L,with_gres", [(1, 32, 256, 3, False), (130, 32, 100, 2, True), (257, 16, 64, 3, False),
                                               (300, 64, 1024, 2, True), (64, 8, 20, 4, False)])
def test_gumbel_backward_vs_oracle(B, D, K, L, with_gres):
    g = torch.Generator().manual_seed(B + D + K)
    x = torch.randn(B, D, generator=g)
    cbs = synthetic_codebooks(torch.randn(max(4 * K, 512), D, generator=g), L, K, seed=3)
    u = torch.rand(L, B, K, generator=g)
    g_emb = torch.randn(L, B, D, generator=g)
    g_loss = torch.rand(B, generator=g)
    g_res = torch.randn(L, B, D, generator=g) if with_gres else None
    T, beta = 0.7, 0.25
    ref, gx_ref, gcb_ref = O.rq_forward_backward(x, cbs, O.MODE_GUMBEL_SOFTMAX, beta, T, g_emb, g_loss,
                                                 uniforms=u, g_res=g_res)
    xd = x.to(DEV).requires_grad_(True)
    cd = torch.stack(cbs).to(DEV).requires_grad_(True)
    emb, res, ids, loss = RF.residual_quantize(xd, cd, RF.MODE_GUMBEL_SOFTMAX, beta, T, RF.DIST_L2, u.to(DEV))
    obj = (emb * g_emb.to(DEV)).sum() + (loss * g_loss.to(DEV)).sum()
    if g_res is not None:
        obj = obj + (res * g_res.to(DEV)).sum()
    obj.backward()
    assert_close(emb.detach().cpu(), ref.embeddings.detach(), "emb")
    assert_close(xd.grad.cpu(), gx_ref, "grad x")
    assert_close(cd.grad.cpu(), gcb_ref, "grad codebooks")


def _philox_uniforms(L, B, K, seed, offset):
    """Host replica of the kernels' noise: Philox4x32-10, counter (row_lo, row_hi, code >> 2, level),
    key (seed ^ offset) split in two words, 24-bit uniforms (csrc/rq_common.cuh)."""
    import numpy as np
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    key = (seed ^ offset) & 0xFFFFFFFFFFFFFFFF
    k0 = np.uint64(key & 0xFFFFFFFF)
    k1 = np.uint64(key >> 32)
    rows = np.arange(B, dtype=np.uint64)
    out = np.zeros((L, B, K), dtype=np.float32)
    mask = np.uint64(0xFFFFFFFF)
    for l in range(L):
        for grp in range((K + 3) // 4):
            c = [rows & mask, rows >> np.uint64(32), np.full(B, grp, np.uint64), np.full(B, l, np.uint64)]
            ka, kb = k0, k1
            for _ in range(10):
                p0, p1 = M0 * c[0], M1 * c[2]
                hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask, p1 >> np.uint64(32), p1 & mask
                c = [hi1 ^ c[1] ^ ka, lo1, hi0 ^ c[3] ^ kb, lo0]
                ka = (ka + np.uint64(0x9E3779B9)) & mask
                kb = (kb + np.uint64(0xBB67AE85)) & mask
            for q in range(4):
                if grp * 4 + q < K:
                    out[l, :, grp * 4 + q] = (c[q] >> np.uint64(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return torch.from_numpy(out)


def test_gumbel_philox_noise_is_reproduced_by_the_backward():
    """In-kernel Philox noise: forward and backward with (seed, offset) must equal forward and backward fed
    the same uniforms explicitly (computed on the host)."""
    B, D, K, L = 70, 32, 50, 2
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, D, generator=g).to(DEV)
    cbs = torch.stack(synthetic_codebooks(torch.randn(512, D, generator=g), L, K, seed=1)).to(DEV)
    g_emb = torch.randn(L, B, D, generator=g).to(DEV)
    g_loss = torch.rand(B, generator=g).to(DEV)
    seed, offset, T = 1234567891011, 7, 0.5
    u = _philox_uniforms(L, B, K, seed, offset).to(DEV)
    e1, r1, i1, l1 = RF.quantize_forward(x, cbs, RF.MODE_GUMBEL_SOFTMAX, 0.25, T, seed=seed, offset=offset)
    e2, r2, i2, l2 = RF.quantize_forward(x, cbs, RF.MODE_GUMBEL_SOFTMAX, 0.25, T, gumbel_uniform=u)
    assert_close(e1.cpu(), e2.cpu(), "philox forward emb", tol=1e-5)
    assert_close(l1.cpu(), l2.cpu(), "philox forward loss", tol=1e-5)
    gx1, gc1 = RF.quantize_backward(g_emb, None, g_loss, r1, i1, cbs, RF.MODE_GUMBEL_SOFTMAX, 0.25, T,
                                    seed=seed, offset=offset)
    gx2, gc2 = RF.quantize_backward(g_emb, None, g_loss, r2, i2, cbs, RF.MODE_GUMBEL_SOFTMAX, 0.25, T, gumbel_uniform=u)
    assert_close(gx1.cpu(), gx2.cpu(), "philox backward g_x", tol=1e-4)
    assert_close(gc1.cpu(), gc2.cpu(), "philox backward g_codebooks", tol=1e-4)


# ---- tensor-core path (rq_gumbel_tc.cu) against the thread-per-row family, same C ABI, impl switch -----------------
@pytest.mark.parametrize("B,D,K,L", [(5000, 64, 1024, 4), (4097, 32, 256, 3), (700, 64, 300, 2), (3000, 32, 1000, 2)])
@pytest.mark.parametrize("noise", ["given", "philox"])
def test_gumbel_forward_tensor_core_vs_simt(B, D, K, L, noise):
    g = torch.Generator().manual_seed(B + K)
    x = torch.randn(B, D, generator=g)
    cbs = torch.stack(synthetic_codebooks(torch.randn(max(4 * K, 2048), D, generator=g), L, K, seed=3))
    gu = torch.rand(L, B, K, generator=g).to(DEV) if noise == "given" else None
    xd, cd = x.to(DEV), cbs.to(DEV)
    kw = dict(gumbel_uniform=gu, seed=11, offset=5)
    e1, r1, i1, l1 = RF.quantize_forward(xd, cd, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, impl=RF.IMPL_SIMT, **kw)
    e2, r2, i2, l2 = RF.quantize_forward(xd, cd, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, impl=RF.IMPL_TCGEN05, **kw)
    torch.cuda.synchronize()
    # ids: both are exact argmins of the same fp32 distances, up to the summation order of |r|^2 + |c|^2 - 2 r.c
    assert float((i1 != i2).float().mean()) < 1e-3
    assert_close(e2, e1, "emb", tol=2e-4)
    assert_close(r2, r1, "res", tol=2e-4)
    assert_close(l2, l1, "loss", tol=2e-4)
    # without the outputs a caller may not want
    _, _, i3, _ = RF.quantize_forward(xd, cd, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, impl=RF.IMPL_TCGEN05,
                                      want_embeddings=False, want_residuals=False, want_loss=False, **kw)
    assert torch.equal(i3, i2)


@pytest.mark.parametrize("B,D,K,L", [(5000, 64, 1024, 4), (4097, 32, 256, 3), (4100, 64, 300, 2), (4500, 32, 1000, 2),
                                     # K <= 64: GEMM + streaming Jacobian pass instead of the fused epilogue; K = 70, 257:
                                     # ragged column tails of the fused epilogue and of the MN-major codebook-gradient GEMMs
                                     (4200, 64, 48, 2), (4300, 32, 70, 2), (4400, 64, 257, 1)])
@pytest.mark.parametrize("noise", ["given", "philox"])
def test_gumbel_backward_tensor_core_vs_thread_per_row(B, D, K, L, noise):
    """rq_quantize_backward in Gumbel mode: the GEMM formulation on the tensor cores against the CUDA-core kernel
    (RQ_BWD_THREAD_PER_ROW), same saved residuals, every upstream gradient present."""
    g = torch.Generator().manual_seed(B + K + 1)
    x = torch.randn(B, D, generator=g)
    cbs = torch.stack(synthetic_codebooks(torch.randn(max(4 * K, 2048), D, generator=g), L, K, seed=3))
    gu = torch.rand(L, B, K, generator=g).to(DEV) if noise == "given" else None
    xd, cd = x.to(DEV), cbs.to(DEV)
    g_emb = torch.randn(L, B, D, generator=g).to(DEV)
    g_res = (torch.randn(L, B, D, generator=g) * 0.1).to(DEV)
    g_loss = torch.rand(B, generator=g).to(DEV)
    _, res, ids, _ = RF.quantize_forward(xd, cd, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2, gumbel_uniform=gu, seed=11, offset=5,
                                         impl=RF.IMPL_SIMT)
    args = (g_emb, g_res, g_loss, res, ids, cd, RF.MODE_GUMBEL_SOFTMAX, 0.25, 0.2)
    kw = dict(gumbel_uniform=gu, seed=11, offset=5)
    gx1, gc1 = RF.quantize_backward(*args, thread_per_row=True, **kw)
    gx2, gc2 = RF.quantize_backward(*args, **kw)
    torch.cuda.synchronize()
    assert_close(gx2, gx1, "g_x", tol=2e-4)
    assert_close(gc2, gc1, "g_codebooks", tol=2e-4)


@pytest.mark.parametrize("B,D,K,L", [(5000, 64, 1024, 4), (4097, 32, 254, 3)])
def test_gumbel_saved_weights_equal_recomputation(B, D, K, L):
    """rq_quantize_forward_ws / rq_quantize_backward_ws: the backward fed with the forward's softmax weights returns
    what the recomputing backward returns (same kernels downstream of W), and the autograd Function uses them."""
    g = torch.Generator().manual_seed(B + K + 2)
    x = torch.randn(B, D, generator=g).to(DEV)
    cbs = torch.stack(synthetic_codebooks(torch.randn(max(4 * K, 2048), D, generator=g), L, K, seed=3)).to(DEV)
    g_emb = torch.randn(L, B, D, generator=g).to(DEV)
    g_loss = torch.rand(B, generator=g).to(DEV)
    M = RF.MODE_GUMBEL_SOFTMAX
    w = RF.gumbel_weights_buffer(x, cbs, M)
    assert w is not None and tuple(w.shape) == (L, B, (K + 3) & ~3)
    emb, res, ids, loss = RF.quantize_forward(x, cbs, M, 0.25, 0.2, seed=4, offset=9, gumbel_weights_out=w)
    assert_close(w[:, :, :K].sum(-1).cpu(), torch.ones(L, B), "rows of W sum to one", tol=1e-5)
    args = (g_emb, None, g_loss, res, ids, cbs, M, 0.25, 0.2)
    gx1, gc1 = RF.quantize_backward(*args, seed=4, offset=9)
    gx2, gc2 = RF.quantize_backward(*args, seed=4, offset=9, gumbel_weights=w)
    assert_close(gx2, gx1, "g_x", tol=1e-5)
    assert_close(gc2, gc1, "g_codebooks", tol=1e-5)
    # through autograd
    xr, cr = x.clone().requires_grad_(True), cbs.clone().requires_grad_(True)
    e, r, i, l = RF.residual_quantize(xr, cr, M, 0.25, 0.2, RF.DIST_L2, None, 4, 9)
    ((e * g_emb).sum() + (l * g_loss).sum()).backward()
    assert_close(xr.grad, gx1, "autograd g_x", tol=1e-5)
    assert_close(cr.grad, gc1, "autograd g_codebooks", tol=1e-5)
    assert RF.gumbel_weights_buffer(x[:100], cbs, M) is None            # small batches: thread-per-row kernels, nothing saved
