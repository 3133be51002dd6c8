"""Losses at the edge of the path (reference: modules/loss.py).

QuantizeLoss is computed INSIDE the fused kernel (per-row (1+beta)*||x-e||^2 with the two stop-
gradient halves handled by the fused backward); the module is kept so `Quantize.quantize_loss`
exists with the same attribute (`commitment_weight`) the reference exposes."""
from torch import Tensor, nn


class ReconstructionLoss(nn.Module):
    def forward(self, x_hat: Tensor, x: Tensor) -> Tensor:
        return ((x_hat - x) ** 2).sum(axis=-1)


class CategoricalReconstuctionLoss(nn.Module):
    def __init__(self, n_cat_feats: int) -> None:
        super().__init__()
        self.reconstruction_loss = ReconstructionLoss()
        self.n_cat_feats = n_cat_feats

    def forward(self, x_hat: Tensor, x: Tensor) -> Tensor:
        n = self.n_cat_feats
        out = self.reconstruction_loss(x_hat[:, :-n], x[:, :-n])
        if n > 0:
            out = out + nn.functional.binary_cross_entropy_with_logits(
                x_hat[:, -n:], x[:, -n:], reduction="none").sum(axis=-1)
        return out


class QuantizeLoss(nn.Module):
    def __init__(self, commitment_weight: float = 1.0) -> None:
        super().__init__()
        self.commitment_weight = commitment_weight

    def forward(self, query: Tensor, value: Tensor) -> Tensor:
        emb_loss = ((query.detach() - value) ** 2).sum(axis=[-1])
        query_loss = ((query - value.detach()) ** 2).sum(axis=[-1])
        return emb_loss + self.commitment_weight * query_loss
