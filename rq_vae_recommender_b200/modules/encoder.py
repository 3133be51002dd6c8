"""Bias-free SiLU MLP (reference: modules/encoder.py:7-36).  Adjacent to the path: its GEMMs stay
on cuBLAS through torch; state-dict keys (`mlp.{0,2,4,6}.weight`) are the reference's."""
from typing import List

from torch import Tensor, nn

from .normalize import L2NormalizationLayer


class MLP(nn.Module):
    def __init__(self, input_dim: int, hidden_dims: List[int], out_dim: int, dropout: float = 0.0,
                 normalize: bool = False) -> None:
        super().__init__()
        self.input_dim, self.hidden_dims, self.out_dim, self.dropout = input_dim, hidden_dims, out_dim, dropout
        dims = [input_dim] + list(hidden_dims) + [out_dim]
        self.mlp = nn.Sequential()
        for i in range(len(dims) - 1):
            self.mlp.append(nn.Linear(dims[i], dims[i + 1], bias=False))
            if i != len(dims) - 2:
                self.mlp.append(nn.SiLU())
                if dropout != 0:
                    self.mlp.append(nn.Dropout(dropout))
        self.mlp.append(L2NormalizationLayer() if normalize else nn.Identity())

    def forward(self, x: Tensor) -> Tensor:
        assert x.shape[-1] == self.input_dim, f"Invalid input dim: Expected {self.input_dim}, found {x.shape[-1]}"
        return self.mlp(x)
