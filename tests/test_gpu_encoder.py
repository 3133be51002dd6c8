"""GPU: the encoder-layer kernel (csrc/rq_linear_tf32.cu; reference MLP modules/encoder.py:17-32) against float64.

Two kinds of check.  (1) Layout: with operands whose fp32 values already fit TF32's 10-bit significand the tensor-core
product is exact up to fp32 accumulation, so the result must agree with float64 to 2e-6 |x||w| — any indexing, swizzle
or zero-fill mistake shows up at full size.  (2) Precision: with ordinary fp32 operands the error must stay inside the
TF32 bound 2^-10 |x||w| (tolerance written below), i.e. the class of error the reference accepts when it sets
torch.set_float32_matmul_precision('high') (modules/rqvae.py:19)."""
import pytest
import torch

from rq_vae_recommender_b200 import _lib
from rq_vae_recommender_b200 import functional as RF

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _tf32_exact(t):
    """round to nearest onto the TF32 grid (10 explicit significand bits)"""
    b = t.contiguous().view(torch.int32)
    b = (b + 0x1000) & ~0x1FFF
    return b.view(torch.float32)


def _ref(x, w, act):
    y = x.double() @ w.double().t()
    if act == RF.ACT_SILU:
        y = y * torch.sigmoid(y)
    elif act == RF.ACT_L2NORM:
        y = y / y.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    return y


SHAPES = [(1000, 768, 512), (4096, 512, 256), (333, 256, 128), (5000, 128, 32), (129, 128, 64), (1, 768, 512),
          (257, 100, 24), (2048, 36, 300), (640, 128, 16),
          (8192, 768, 512), (5000, 256, 768), (4097, 64, 1024), (70000, 128, 512)]


@pytest.mark.parametrize("rows,n_in,n_out", SHAPES)
@pytest.mark.parametrize("act", [RF.ACT_NONE, RF.ACT_SILU])
def test_encoder_layer_layout_exact_on_tf32_grid(rows, n_in, n_out, act):
    g = torch.Generator(device=DEV).manual_seed(rows * 7 + n_in + n_out)
    x = _tf32_exact(torch.randn(rows, n_in, device=DEV, generator=g))
    w = _tf32_exact(torch.randn(n_out, n_in, device=DEV, generator=g) / n_in ** 0.5)
    y = RF.encoder_layer(x, w, act)
    ref = _ref(x, w, act)
    bound = x.double().norm(dim=1, keepdim=True) * w.double().norm(dim=1)[None, :]
    err = ((y.double() - ref).abs() / bound).max().item()
    assert err < 2e-6, err
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0


@pytest.mark.parametrize("rows,n_in,n_out,act", [(3000, 768, 512, RF.ACT_SILU), (3000, 128, 32, RF.ACT_NONE),
                                                 (999, 128, 64, RF.ACT_L2NORM), (4100, 128, 32, RF.ACT_L2NORM),
                                                 (500, 64, 48, RF.ACT_L2NORM), (300, 256, 8, RF.ACT_L2NORM)])
def test_encoder_layer_tf32_precision(rows, n_in, n_out, act):
    g = torch.Generator(device=DEV).manual_seed(rows + n_in + n_out + act)
    x = torch.randn(rows, n_in, device=DEV, generator=g)
    w = torch.randn(n_out, n_in, device=DEV, generator=g) / n_in ** 0.5
    y = RF.encoder_layer(x, w, act)
    ref = _ref(x, w, act)
    if act == RF.ACT_L2NORM:
        assert torch.allclose(y.norm(dim=1), torch.ones(rows, device=DEV), atol=1e-5)
        err = (y.double() - ref).abs().max().item()
        assert err < 4e-3, err                             # unit rows: 2^-10-grade products, a few of them per entry
    else:
        bound = x.double().norm(dim=1, keepdim=True) * w.double().norm(dim=1)[None, :]
        err = ((y.double() - ref).abs() / bound).max().item()
        assert err < 2.0 ** -10, err
    assert torch.isfinite(y).all()


@pytest.mark.parametrize("rows,n_in,n_out", [(8192, 768, 512), (5000, 256, 768), (4097, 64, 1024), (70000, 128, 512),
                                              (513, 96, 512)])
def test_encoder_layer_cta_pair_schedule(rows, n_in, n_out):
    """ACT_CTA_PAIR: clusters of two CTAs, cta_group::2 MMAs, each CTA staging half of the weight tile - same result as
    the default schedule bit for bit (same products, same accumulation order), ragged row tails included."""
    g = torch.Generator(device=DEV).manual_seed(rows + n_in)
    x = torch.randn(rows, n_in, device=DEV, generator=g)
    w = torch.randn(n_out, n_in, device=DEV, generator=g) / n_in ** 0.5
    for act in (RF.ACT_NONE, RF.ACT_SILU | RF.ACT_TF32_OUT):
        y1 = RF.encoder_layer(x, w, act)
        y2 = RF.encoder_layer(x, w, act | RF.ACT_CTA_PAIR)
        assert torch.equal(y1, y2)
    xe, we = _tf32_exact(x), _tf32_exact(w)
    y = RF.encoder_layer(xe, we, RF.ACT_CTA_PAIR)
    bound = xe.double().norm(dim=1, keepdim=True) * we.double().norm(dim=1)[None, :]
    assert ((y.double() - _ref(xe, we, RF.ACT_NONE)).abs() / bound).max().item() < 2e-6
    assert _lib.lib().rq_debug_tc_timeout_flag() == 0


def test_encoder_layer_rejects_bad_shapes():
    x = torch.randn(16, 30, device=DEV)
    w = torch.randn(8, 30, device=DEV)
    with pytest.raises(_lib.RqEngineError):
        RF.encoder_layer(x, w)                             # in_features not a multiple of 4
    x = torch.randn(16, 128, device=DEV)
    w = torch.randn(128, 128, device=DEV)
    with pytest.raises(_lib.RqEngineError):
        RF.encoder_layer(x, w, RF.ACT_L2NORM)              # the fused normalisation holds a whole row per thread
    assert not RF.encoder_layer_supported(16, 128, 128, RF.ACT_L2NORM)
    assert RF.encoder_layer_supported(16, 128, 64, RF.ACT_L2NORM)


@pytest.mark.parametrize("d,normalize", [(32, False), (64, True)])
def test_mlp_inference_runs_on_engine_and_matches_torch_tf32(d, normalize):
    """The reference encoder shape (768-512-256-128-d, modules/encoder.py) in inference: the mirror's MLP takes the engine
    path when TF32 matmuls are allowed, its result is as close to float64 as torch's own TF32 chain (within 1.5x: the
    tensor core truncates the raw fp32 input of the first layer, everything after it is rounded to nearest), and with
    matmul precision 'highest' (or autograd) it stays on torch bit for bit."""
    from rq_vae_recommender_b200.modules.encoder import MLP
    torch.manual_seed(3)
    mlp = MLP(768, [512, 256, 128], d, normalize=normalize).to(DEV).eval()
    x = torch.nn.functional.normalize(torch.randn(10000, 768, device=DEV), dim=-1)
    prev = torch.get_float32_matmul_precision()
    try:
        with torch.no_grad():
            ref = mlp.double()(x.double())
            mlp.float()
            torch.set_float32_matmul_precision("highest")
            assert mlp._engine_plan(x) is None
            y_fp32 = mlp(x)
            assert torch.equal(y_fp32, mlp.mlp(x))
            torch.set_float32_matmul_precision("high")
            assert mlp._engine_plan(x) is not None
            before = _lib.lib().rq_launch_count()
            y = mlp(x)
            assert _lib.lib().rq_launch_count() - before == 4          # one kernel per Linear, activations fused
            y_torch_tf32 = mlp.mlp(x)
        with torch.enable_grad():
            assert mlp._engine_plan(x) is None                          # parameters require grad: autograd path
    finally:
        torch.set_float32_matmul_precision(prev)
    rel = ((y.double() - ref).norm() / ref.norm()).item()
    rel_torch = ((y_torch_tf32.double() - ref).norm() / ref.norm()).item()
    assert rel < 1e-3 and rel <= 1.5 * rel_torch + 1e-5, (rel, rel_torch)
    assert y.shape == y_fp32.shape and torch.isfinite(y).all()
    if normalize:
        assert torch.allclose(y.norm(dim=1), torch.ones(len(y), device=DEV), atol=1e-5)


def test_tf32_out_flag_rounds_to_grid():
    x = torch.randn(1000, 128, device=DEV)
    w = torch.randn(256, 128, device=DEV) / 11.3
    y = RF.encoder_layer(x, w, RF.ACT_SILU | RF.ACT_TF32_OUT)
    assert torch.equal(y, _tf32_exact(y)) and (y.view(torch.int32) & 0x1FFF).eq(0).all()
    y0 = RF.encoder_layer(x, w, RF.ACT_SILU)
    assert torch.equal(y, _tf32_exact(y0))


def test_encoder_chain_captures_into_a_cuda_graph():
    """The four layers of the reference encoder captured into one CUDA graph on a fresh stream (nothing on the launch
    path allocates or synchronises) and replayed on new inputs: same bits as the eager calls."""
    torch.manual_seed(11)
    dims = [768, 512, 256, 128, 32]
    ws = [_tf32_exact(torch.randn(dims[i + 1], dims[i], device=DEV) / dims[i] ** 0.5) for i in range(4)]
    x = torch.randn(6000, 768, device=DEV)
    bufs = [torch.empty(6000, d, device=DEV) for d in dims[1:]]

    def chain():
        h = x
        for i, w in enumerate(ws):
            act = (RF.ACT_SILU | RF.ACT_TF32_OUT) if i < 3 else RF.ACT_L2NORM
            h = RF.encoder_layer(h, w, act, out=bufs[i])
        return h

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        chain()                                             # warm-up outside the capture (function attributes)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        chain()
    x.copy_(torch.randn(6000, 768, device=DEV))
    g.replay()
    got = bufs[3].clone()
    want = chain().clone()
    assert torch.equal(got, want)
    assert torch.allclose(got.norm(dim=1), torch.ones(6000, device=DEV), atol=1e-5)
