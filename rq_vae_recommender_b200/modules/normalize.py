"""L2 normalisation helpers (reference: modules/normalize.py:7-18).  Stay on PyTorch: they act on
the K x D effective codebook and the encoder output, both negligible next to the quantise kernel."""
import torch.nn.functional as F
from torch import Tensor, nn


def l2norm(x: Tensor, dim: int = -1, eps: float = 1e-12) -> Tensor:
    return F.normalize(x, p=2, dim=dim, eps=eps)


class L2NormalizationLayer(nn.Module):
    def __init__(self, dim: int = -1, eps: float = 1e-12) -> None:
        super().__init__()
        self.dim = dim
        self.eps = eps

    def forward(self, x: Tensor) -> Tensor:
        return l2norm(x, dim=self.dim, eps=self.eps)
