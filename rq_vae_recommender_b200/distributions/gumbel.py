"""Gumbel noise helpers (reference: distributions/gumbel.py:7-17).

In the engine the Gumbel-softmax itself is fused into the quantise kernel; these torch versions
exist for API parity and for drawing the uniform tensor in noise-parity mode."""
from typing import Tuple

import torch
import torch.nn.functional as F
from torch import Tensor


def sample_gumbel(shape: Tuple, device: torch.device, eps: float = 1e-20) -> Tensor:
    u = torch.rand(shape, device=device)
    return -torch.log(-torch.log(u + eps) + eps)


def gumbel_softmax_sample(logits: Tensor, temperature: float, device: torch.device) -> Tensor:
    return F.softmax((logits + sample_gumbel(logits.shape, device)) / temperature, dim=-1)
