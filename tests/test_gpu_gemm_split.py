"""GPU: the fp32-grade tensor-core GEMM of the Gumbel-softmax path (csrc/rq_gemm_split.cu: three bf16 planes per fp32
value, six tcgen05 products) against float64, at the three shapes the path uses it in — [B, D] x [K, D]^T, [B, K] x
[D, K]^T and the split-K [K, B] x [D, B]^T — with ragged sizes, strides, alpha and atomic accumulation."""
import pytest
import torch

from rq_vae_recommender_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _gemm(A, B, C, alpha=1.0, atomic=0, k_splits=1):
    M, Kd = A.shape
    N = B.shape[0]
    rc = _lib.lib().rq_debug_split_gemm(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, Kd, A.stride(0), B.stride(0),
                                        C.stride(0), float(alpha), int(atomic), int(k_splits),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _lib.last_error() if hasattr(_lib, "last_error") else rc
    return C


@pytest.mark.parametrize("M,N,Kd", [(1000, 300, 64), (128, 256, 64), (777, 64, 1000), (4097, 1024, 32), (130, 40, 200),
                                      (1, 1, 4)])
def test_split_gemm_matches_float64(M, N, Kd):
    g = torch.Generator(device=DEV).manual_seed(M + N + Kd)
    A = torch.randn(M, Kd, device=DEV, generator=g) * 3.0
    B = torch.randn(N, Kd, device=DEV, generator=g)
    A[::7] *= 1e-3                                   # rows of very different magnitude: the split is per element
    C = torch.full((M, N), float("nan"), device=DEV)
    _gemm(A, B, C, alpha=-2.0)
    torch.cuda.synchronize()
    want = -2.0 * (A.double() @ B.double().T)
    scale = (A.double().norm(dim=1)[:, None] * B.double().norm(dim=1)[None, :]).clamp_min(1e-30) * 2.0
    err = ((C.double() - want).abs() / scale).max()
    # the TMEM accumulation truncates (2^-23 per MMA on the running sum): a few 1e-7 of |a||b|, where a bf16 product is
    # 8e-3 and a two-term split 3e-5
    assert float(err) < 1.5e-6, f"relative to |a||b|: {float(err):.2e}"


def test_split_gemm_strided_operands_and_split_k_accumulation():
    g = torch.Generator(device=DEV).manual_seed(9)
    Kc, D, Bn = 1024, 64, 5000
    Wt = torch.randn(Kc, Bn + 24, device=DEV, generator=g)[:, :Bn]           # row stride 5024
    det = torch.randn(D, Bn, device=DEV, generator=g)
    C = torch.randn(Kc, D + 8, device=DEV, generator=g)[:, :D]              # live buffer, row stride 72
    C0 = C.clone()
    _gemm(Wt, det, C, alpha=0.5, atomic=1, k_splits=8)
    torch.cuda.synchronize()
    want = C0.double() + 0.5 * (Wt.double() @ det.double().T)
    scale = (Wt.double().norm(dim=1)[:, None] * det.double().norm(dim=1)[None, :]) * 0.5
    assert float(((C.double() - want).abs() / scale).max()) < 1.5e-6


def _gemm_tn(At, Bt, C, alpha=1.0, atomic=0, k_splits=1):
    Kd, M = At.shape
    N = Bt.shape[1]
    rc = _lib.lib().rq_debug_split_gemm_tn(At.data_ptr(), Bt.data_ptr(), C.data_ptr(), M, N, Kd, At.stride(0), Bt.stride(0),
                                           C.stride(0), float(alpha), int(atomic), int(k_splits),
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0, rc
    return C


@pytest.mark.parametrize("M,N,Kd,splits", [(1024, 64, 65536, 37), (256, 32, 4096, 1), (1000, 64, 777, 3), (130, 20, 200, 2),
                                           (128, 64, 64, 1), (4, 4, 5, 1)])
def test_split_gemm_transposed_operands_match_float64(M, N, Kd, splits):
    """C = At^T Bt with At [Kd, M], Bt [Kd, N] as they lie in memory (the Gumbel backward's W^T de over the batch): staged
    MN-major, no transposed copies.  fp32-grade (six products) and the two-plane form, split-K with atomics."""
    g = torch.Generator(device=DEV).manual_seed(M + N + Kd)
    ldm, ldn = (M + 3) // 4 * 4 + 8, (N + 3) // 4 * 4
    At = (torch.randn(Kd, ldm, device=DEV, generator=g) * 2.0)[:, :M]
    Bt = torch.randn(Kd, ldn, device=DEV, generator=g)[:, :N]
    ref = At.double().t() @ Bt.double()
    bound = At.double().norm(dim=0)[:, None] * Bt.double().norm(dim=0)[None, :]
    C = torch.zeros(M, N, device=DEV)
    _gemm_tn(At, Bt, C, alpha=1.0, atomic=1, k_splits=splits)
    err = ((C.double() - ref).abs() / bound).max().item()
    assert err < 2.5e-6, err        # fp32 accumulation (truncating) over up to 4,096 products in one split
    C3 = torch.zeros(M, N, device=DEV)
    _gemm_tn(At, Bt, C3, alpha=-2.0, atomic=1, k_splits=-splits)
    err3 = ((C3.double() + 2.0 * ref).abs() / bound).max().item()
    assert err3 < 2.0 ** -14, err3
