"""GPU: the tcgen05 family against the exact SIMT family (same C ABI, impl switch) and the oracle."""
import pytest
import torch

import rq_oracle as O
import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import _lib
from rq_vae_recommender_b200 import functional as RF
from parity_utils import assert_close, check_ids, matching_rows, synthetic_codebooks

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(B, K, L, seed, scale=1.0, D=32):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, D, generator=g) * scale
    cbs = synthetic_codebooks(torch.randn(max(4 * K, 2048), D, generator=g) * scale, L, K, seed=seed + 1)
    return x, cbs


def test_tc_is_selected_for_the_headline_shape():
    assert _lib.lib().rq_tcgen05_supported(32, 256, 3, RF.MODE_ROTATION_TRICK, RF.DIST_L2) == 1
    assert _lib.lib().rq_tcgen05_supported(64, 256, 3, RF.MODE_EVAL, RF.DIST_L2) == 1      # ml32m shape
    assert _lib.lib().rq_tcgen05_supported(64, 256, 4, RF.MODE_EVAL, RF.DIST_L2) == 1      # level-major composition
    assert _lib.lib().rq_tcgen05_supported(64, 1024, 4, RF.MODE_ROTATION_TRICK, RF.DIST_L2) == 1   # config 5 shape
    assert _lib.lib().rq_tcgen05_supported(64, 1024, 4, RF.MODE_GUMBEL_SOFTMAX, RF.DIST_L2) == 1   # rq_gumbel_tc.cu
    assert _lib.lib().rq_tcgen05_supported(64, 1024, 4, RF.MODE_GUMBEL_SOFTMAX, RF.DIST_COSINE) == 0
    assert _lib.lib().rq_tcgen05_supported(32, 256, 3, RF.MODE_EVAL, RF.DIST_COSINE) == 0


def test_tc_accumulator_error_bound():
    """Raw tensor-core accumulator (||c||^2 - 2 x.c) of tile 0 / level 0 vs float64: the candidate
    threshold in the kernel is only sound if this error stays inside the stated bound."""
    x, cbs = _inputs(128, 256, 1, seed=11)
    buf = torch.zeros(128, 256, device=DEV)
    _lib.lib().rq_debug_tc_set_accumulator_dump(buf.data_ptr())
    try:
        RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), RF.MODE_EVAL, impl=RF.IMPL_TCGEN05)
        torch.cuda.synchronize()
    finally:
        _lib.lib().rq_debug_tc_set_accumulator_dump(None)
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
    c = cbs[0].double()
    want = (c ** 2).sum(1)[None, :] - 2.0 * x.double() @ c.T
    err = (buf.cpu().double() - want).abs()
    xn = x.double().norm(dim=1, keepdim=True)
    cn = c.norm(dim=1)[None, :]
    bound = 0.5 * (1e-4 * xn * cn.max() + 3e-6 * cn.max() ** 2)       # the kernel uses 2E as threshold
    assert bool((err <= bound).all()), f"max err {float(err.max()):.3e}, bound min {float(bound.min()):.3e}"
    # and it is far tighter than a single bf16 pass would be
    assert float((err / (xn * cn)).max()) < 4e-5


@pytest.mark.parametrize("B,K,L", [(128, 256, 1), (1, 256, 3), (300, 256, 3), (4097, 200, 2), (65536, 256, 3),
                                    (70001, 256, 4), (1000, 16, 3)])
def test_tc_ids_bit_identical_to_simt(B, K, L):
    x, cbs = _inputs(B, K, L, seed=B % 1000 + K)
    xd, cd = x.to(DEV), torch.stack(cbs).to(DEV)
    for mode in (RF.MODE_EVAL, RF.MODE_ROTATION_TRICK, RF.MODE_STE):
        e1, r1, i1, l1 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_SIMT)
        e2, r2, i2, l2 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_TCGEN05)
        torch.cuda.synchronize()
        assert _lib.lib().rq_debug_tc_timeout_flag() == 0
        n_bad = int((i1 != i2).sum())
        assert n_bad == 0, f"mode {mode}: {n_bad} ids differ between tcgen05 and SIMT families"
        assert_close(e2, e1, "emb", tol=1e-6)
        assert_close(r2, r1, "res", tol=1e-6)
        assert_close(l2, l1, "loss", tol=1e-6)
    _, _, i3, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                      want_loss=False, impl=RF.IMPL_TCGEN05)
    _, _, i1, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    assert torch.equal(i1, i3)


@pytest.mark.parametrize("scale", [1e-3, 1.0, 50.0])
def test_tc_vs_oracle_scales_and_near_ties(scale):
    """Different magnitudes, duplicated codewords (exact ties -> first index), and rows sitting
    almost on the bisector of two codes (the re-scoring path)."""
    B, K, L = 6000, 256, 3
    x, cbs = _inputs(B, K, L, seed=3, scale=scale)
    cbs[0][17] = cbs[0][5]                          # exact duplicate
    g2 = torch.Generator().manual_seed(77)
    x[:200] = 0.5 * (cbs[0][torch.arange(200) % K] + cbs[0][(torch.arange(200) + 1) % K]) \
        + 1e-7 * scale * torch.randn(200, 32, generator=g2)
    x[200:260] = cbs[0][5] + 1e-4 * scale * torch.randn(60, 32, generator=g2)
    ref = O.rq_forward(x, cbs, O.MODE_EVAL)
    emb, res, ids, loss = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), RF.MODE_EVAL, impl=RF.IMPL_TCGEN05)
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
    assert int((ids[0] == 17).sum()) == 0           # first index wins
    check_ids(ids, ref, cbs, f"scale {scale}")
    ok = matching_rows(ids, ref)
    assert_close(emb.cpu()[:, ok], ref.embeddings[:, ok], "emb")
    assert_close(loss.cpu()[ok], ref.quantize_loss[ok], "loss")
    _, _, ids_s, _ = RF.quantize_forward(x.to(DEV), torch.stack(cbs).to(DEV), RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    assert torch.equal(ids, ids_s)


# ---- D = 64 (ml32m shape): scaled-fp16 operands, rq_quantize_tc64.cu ------------------------------------------
@pytest.mark.parametrize("B,K,L", [(128, 256, 1), (1, 256, 3), (300, 256, 3), (4097, 200, 2), (40000, 256, 3),
                                    (1000, 16, 3), (19073, 255, 1)])
def test_tc64_ids_bit_identical_to_simt(B, K, L):
    x, cbs = _inputs(B, K, L, seed=B % 1000 + K + 64, D=64)
    xd, cd = x.to(DEV), torch.stack(cbs).to(DEV)
    for mode in (RF.MODE_EVAL, RF.MODE_ROTATION_TRICK, RF.MODE_STE):
        e1, r1, i1, l1 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_SIMT)
        e2, r2, i2, l2 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_TCGEN05)
        torch.cuda.synchronize()
        assert _lib.lib().rq_debug_tc_timeout_flag() == 0
        n_bad = int((i1 != i2).sum())
        assert n_bad == 0, f"mode {mode}: {n_bad} ids differ between tcgen05 (d=64) and SIMT families"
        assert_close(e2, e1, "emb", tol=1e-6)
        assert_close(r2, r1, "res", tol=1e-6)
        assert_close(l2, l1, "loss", tol=1e-6)
    _, _, i3, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                      want_loss=False, impl=RF.IMPL_TCGEN05)
    _, _, i1, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    assert torch.equal(i1, i3)


@pytest.mark.parametrize("scale", [1e-4, 1.0, 300.0])
def test_tc64_vs_oracle_scales_ties_and_outliers(scale):
    """fp16 operands carry a per-level power-of-two scale: magnitudes far from 1, exact duplicate
    codewords, rows on a bisector, rows that dwarf the codebook (skip the tensor-core score and are
    re-scored against all codes) and all-zero rows must all give the SIMT / oracle ids."""
    B, K, L, D = 6000, 256, 3, 64
    x, cbs = _inputs(B, K, L, seed=5, scale=scale, D=D)
    cbs[0][17] = cbs[0][5]
    g2 = torch.Generator().manual_seed(78)
    x[:200] = 0.5 * (cbs[0][torch.arange(200) % K] + cbs[0][(torch.arange(200) + 1) % K]) \
        + 1e-7 * scale * torch.randn(200, D, generator=g2)
    x[200:260] = cbs[0][5] + 1e-4 * scale * torch.randn(60, D, generator=g2)
    x[300:340] *= 3e5                                # |t r| > 2^14: the fp16 overflow guard
    x[340:350] *= 1e-6                               # far below the codebook scale
    x[350:360] = 0.0
    xd, cd = x.to(DEV), torch.stack(cbs).to(DEV)
    _, _, ids_s, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    emb, res, ids, loss = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_TCGEN05)
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
    assert torch.equal(ids, ids_s)
    assert int((ids[0] == 17).sum()) == 0
    keep = torch.ones(B, dtype=torch.bool)
    keep[300:340] = False                            # the oracle's own fp32 cancellation differs at 3e5x
    ref = O.rq_forward(x[keep], cbs, O.MODE_EVAL)
    check_ids(ids[:, keep.to(DEV)], ref, cbs, f"d64 scale {scale}")
    ok = matching_rows(ids[:, keep.to(DEV)], ref)
    assert_close(emb.cpu()[:, keep][:, ok], ref.embeddings[:, ok], "emb")
    assert_close(loss.cpu()[keep][ok], ref.quantize_loss[ok], "loss")


def test_tc64_codebook_scale_extremes():
    """Level scales span 2^-100..2^100 in the kernel; codebooks of very different magnitude per level
    (and an all-zero level) keep ids identical to the SIMT family."""
    B, K, D = 3000, 256, 64
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, D, generator=g)
    cbs = torch.stack([torch.randn(K, D, generator=g) * 1e3, torch.randn(K, D, generator=g) * 1e-5,
                       torch.zeros(K, D)])
    xd, cd = x.to(DEV), cbs.to(DEV)
    _, _, i1, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    _, _, i2, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_TCGEN05)
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
    assert torch.equal(i1, i2)


# ---- D = 64, K > 256 or L > 3: level-major composition over 256-code chunks (rq_quantize_bigk.cu) ------------------
@pytest.mark.parametrize("B,K,L", [(300, 1024, 4), (5000, 1000, 2), (20000, 512, 3), (4097, 256, 4), (1, 1024, 1),
                                    (65536, 1024, 4)])
def test_tc64_bigk_ids_bit_identical_to_simt(B, K, L):
    x, cbs = _inputs(B, K, L, seed=B % 1000 + K + 7, D=64)
    xd, cd = x.to(DEV), torch.stack(cbs).to(DEV)
    for mode in (RF.MODE_EVAL, RF.MODE_ROTATION_TRICK, RF.MODE_STE):
        e1, r1, i1, l1 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_SIMT)
        e2, r2, i2, l2 = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_TCGEN05)
        torch.cuda.synchronize()
        assert _lib.lib().rq_debug_tc_timeout_flag() == 0
        n_bad = int((i1 != i2).sum())
        assert n_bad == 0, f"mode {mode}: {n_bad} ids differ between the chunked tcgen05 path and the SIMT family"
        assert_close(e2, e1, "emb", tol=1e-6)
        assert_close(r2, r1, "res", tol=1e-6)
        assert_close(l2, l1, "loss", tol=1e-6)
    _, _, i3, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, want_embeddings=False, want_residuals=False,
                                      want_loss=False, impl=RF.IMPL_TCGEN05)
    _, _, i1, _ = RF.quantize_forward(xd, cd, RF.MODE_EVAL, impl=RF.IMPL_SIMT)
    assert torch.equal(i1, i3)


@pytest.mark.parametrize("B,K,L", [(70001, 700, 8), (129, 300, 5), (33000, 256, 3)])
def test_tc64_streamed_kernel_shapes(B, K, L):
    """The streamed-chunk kernel at shapes the round-1 tests did not reach: 8 levels, a partial last chunk, an odd
    number of row tiles (the second warpgroup's last tile is empty), and K <= 256 (one chunk per level: since round 2
    every D = 64 tcgen05 launch goes through this kernel)."""
    x, cbs = _inputs(B, K, L, seed=B % 1000 + K + 11, D=64)
    xd, cd = x.to(DEV), torch.stack(cbs).to(DEV)
    e1, r1, i1, l1 = RF.quantize_forward(xd, cd, RF.MODE_ROTATION_TRICK, impl=RF.IMPL_SIMT)
    e2, r2, i2, l2 = RF.quantize_forward(xd, cd, RF.MODE_ROTATION_TRICK, impl=RF.IMPL_TCGEN05)
    torch.cuda.synchronize()
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
    assert int((i1 != i2).sum()) == 0
    assert_close(e2, e1, "emb", tol=1e-6)
    assert_close(r2, r1, "res", tol=1e-6)
    assert_close(l2, l1, "loss", tol=1e-6)


def test_tc64_streamed_kernel_ties_across_chunks():
    """The merge over the chunks of a level: exact duplicates of a code in later chunks (the lowest index must win, as
    in the reference's first-index argmin, quantize.py:121), codes that differ from a code of another chunk in the last
    bits (settled by the exact fp32 distances), rows that sit exactly on a codeword, and a row scale that sends the
    scores of different chunks close together."""
    g = torch.Generator().manual_seed(5)
    D, K, L, B = 64, 1024, 2, 6000
    base = torch.randn(256, D, generator=g)
    cb0 = torch.cat([base,                                        # chunk 0
                     base.clone(),                                # chunk 1: exact duplicates of chunk 0
                     base * (1 + 1e-7) + 1e-8,                    # chunk 2: last-bit neighbours
                     torch.randn(256, D, generator=g)])           # chunk 3: unrelated
    cb0[512 + 7] = base[7]                                        # one exact duplicate inside the near-tie chunk
    cb1 = torch.cat([torch.randn(512, D, generator=g) * 0.3, torch.zeros(256, D), torch.randn(256, D, generator=g) * 0.3])
    cbs = torch.stack([cb0, cb1])
    x = torch.randn(B, D, generator=g)
    x[:256] = base                                                # rows on a codeword (distance 0 in three chunks)
    x[256:512] = 0.5 * (base + cb0[768:1024])                     # bisectors between chunk 0 / 1 / 2 codes and chunk 3
    x[512:600] = 0.0
    xd, cd = x.to(DEV), cbs.to(DEV)
    for mode in (RF.MODE_EVAL, RF.MODE_ROTATION_TRICK):
        _, _, i1, _ = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_SIMT)
        _, _, i2, _ = RF.quantize_forward(xd, cd, mode, impl=RF.IMPL_TCGEN05)
        torch.cuda.synchronize()
        assert _lib.lib().rq_debug_tc_timeout_flag() == 0
        assert int((i1 != i2).sum()) == 0
    # chunk 1 holds bit-identical copies of chunk 0: their exact distances are equal, so the copy can never be chosen
    assert int(((i2[0] >= 256) & (i2[0] < 512)).sum()) == 0


@pytest.mark.parametrize("K,L", [(256, 3), (1024, 4)])
def test_tc64_streamed_kernel_in_cuda_graph(K, L):
    """The D = 64 path allocates its operand-image scratch from the library's stream-ordered pool: captured into a CUDA
    graph (allocation and free become graph nodes) the replays must equal eager launches on the same inputs."""
    g = torch.Generator(device=DEV).manual_seed(K + L)
    D, B = 64, 20000
    cb = torch.stack([torch.randn(K, D, device=DEV, generator=g) * (0.6 ** l) for l in range(L)])
    xs = torch.randn(B, D, device=DEV, generator=g)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        RF.quantize_forward(xs, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = RF.quantize_forward(xs, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2)
    for _ in range(2):
        x2 = torch.randn(B, D, device=DEV, generator=g)
        xs.copy_(x2)
        graph.replay()
        torch.cuda.synchronize()
        ref = RF.quantize_forward(x2, cb, RF.MODE_ROTATION_TRICK, 0.25, 0.2)
        for a, b in zip(out, ref):
            assert torch.equal(a, b)
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0
