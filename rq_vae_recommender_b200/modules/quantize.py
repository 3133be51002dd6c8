"""One vector-quantisation level behind the reference's `Quantize` interface.

Mirrors modules/quantize.py of AdamLTy/RQ-VAE-Recommender: same constructor, attributes,
`forward(x, temperature) -> QuantizeOutput`, enum values and error behaviour.  The arithmetic of
:105-150 (distance, argmin, the three estimators, the loss) runs in ONE CUDA launch of the fused
kernel with L = 1; `RqVae.get_semantic_ids` fuses all levels into one launch instead of calling
this per level.
"""
from enum import Enum
from typing import NamedTuple, Optional

import torch
from torch import Tensor, nn

from .. import functional as RF
from .._gin import constants_from_enum
from ..init.kmeans import kmeans_init_
from .loss import QuantizeLoss
from .normalize import L2NormalizationLayer


@constants_from_enum(module="modules.quantize")
class QuantizeForwardMode(Enum):
    GUMBEL_SOFTMAX = 1
    STE = 2
    ROTATION_TRICK = 3


class QuantizeDistance(Enum):
    L2 = 1
    COSINE = 2


class QuantizeOutput(NamedTuple):
    embeddings: Tensor
    ids: Tensor
    loss: Tensor


def kernel_mode(forward_mode, training: bool) -> int:
    """Enum -> C-ABI mode; the `not self.training` branch (quantize.py:148-150) is mode 0."""
    if not training:
        if not isinstance(forward_mode, QuantizeForwardMode):
            return RF.MODE_EVAL
        return RF.MODE_EVAL
    if forward_mode == QuantizeForwardMode.GUMBEL_SOFTMAX:
        return RF.MODE_GUMBEL_SOFTMAX
    if forward_mode == QuantizeForwardMode.STE:
        return RF.MODE_STE
    if forward_mode == QuantizeForwardMode.ROTATION_TRICK:
        return RF.MODE_ROTATION_TRICK
    raise Exception("Unsupported Quantize forward mode.")           # quantize.py:144


def kernel_distance(distance_mode) -> int:
    if distance_mode == QuantizeDistance.L2:
        return RF.DIST_L2
    if distance_mode == QuantizeDistance.COSINE:
        return RF.DIST_COSINE
    raise Exception("Unsupported Quantize distance mode.")          # quantize.py:119


class GumbelNoise:
    """Where the Gumbel-softmax uniforms come from.

    "philox": drawn inside the kernel (Philox4x32-10 keyed by torch.initial_seed() and a call
    counter) — nothing but the latents is read from HBM.  "torch": `torch.rand((B, K))` per level
    on the module's device, i.e. exactly the draws gumbel.py:9 makes, handed to the kernel — the
    noise-parity mode."""
    source = "philox"
    _calls = 0

    @classmethod
    def next(cls, n_levels: int, B: int, K: int, device):
        # under torch.compile a Python-side call counter would be baked into the graph as a constant (the same noise
        # every step): there the draws are torch.rand nodes of the graph, like the reference's (gumbel.py:9)
        if cls.source == "torch" or torch.compiler.is_compiling():
            return torch.stack([torch.rand((B, K), device=device) for _ in range(n_levels)]), 0, 0
        cls._calls += 1
        return None, int(torch.initial_seed()) & 0xFFFFFFFFFFFFFFFF, cls._calls


class Quantize(nn.Module):
    def __init__(
        self,
        embed_dim: int,
        n_embed: int,
        do_kmeans_init: bool = True,
        codebook_normalize: bool = False,
        sim_vq: bool = False,
        commitment_weight: float = 0.25,
        forward_mode: QuantizeForwardMode = QuantizeForwardMode.GUMBEL_SOFTMAX,
        distance_mode: QuantizeDistance = QuantizeDistance.L2,
    ) -> None:
        super().__init__()
        self.embed_dim = embed_dim
        self.n_embed = n_embed
        self.embedding = nn.Embedding(n_embed, embed_dim)
        self.forward_mode = forward_mode
        self.distance_mode = distance_mode
        self.do_kmeans_init = do_kmeans_init
        self.kmeans_initted = False          # plain attribute, not in the state dict (quantize.py:68)
        self.out_proj = nn.Sequential(
            nn.Linear(embed_dim, embed_dim, bias=False) if sim_vq else nn.Identity(),
            L2NormalizationLayer(dim=-1) if codebook_normalize else nn.Identity(),
        )
        self.quantize_loss = QuantizeLoss(commitment_weight)
        self._init_weights()

    @property
    def weight(self) -> Tensor:
        return self.embedding.weight

    @property
    def device(self) -> torch.device:
        return self.embedding.weight.device

    def _init_weights(self) -> None:
        nn.init.uniform_(self.embedding.weight)                      # U[0,1), quantize.py:86-89

    @torch.no_grad()
    def _kmeans_init(self, x: Tensor) -> None:
        kmeans_init_(self.embedding.weight, x=x)
        self.kmeans_initted = True

    def needs_kmeans_init(self) -> bool:
        return bool(self.do_kmeans_init and not self.kmeans_initted)

    def effective_codebook(self) -> Tensor:
        """`out_proj(embedding.weight)` (quantize.py:105) — stays in PyTorch so that sim_vq and
        codebook_normalize gradients flow by autograd."""
        return self.out_proj(self.embedding.weight)

    def get_item_embeddings(self, item_ids: Tensor) -> Tensor:
        return self.out_proj(self.embedding(item_ids))

    def forward(self, x: Tensor, temperature: float) -> QuantizeOutput:
        assert x.shape[-1] == self.embed_dim                         # quantize.py:100
        if self.needs_kmeans_init():
            self._kmeans_init(x=x)
        distance = kernel_distance(self.distance_mode)
        mode = kernel_mode(self.forward_mode, self.training)
        codebook = self.effective_codebook().unsqueeze(0)            # [1, K, D]
        gu, seed, offset = (None, 0, 0)
        if mode == RF.MODE_GUMBEL_SOFTMAX:
            gu, seed, offset = GumbelNoise.next(1, x.shape[0], self.n_embed, x.device)
        emb, _res, ids, loss = RF.residual_quantize(
            x, codebook, mode, float(self.quantize_loss.commitment_weight), float(temperature), distance,
            gu, seed, offset)
        return QuantizeOutput(embeddings=emb[0], ids=ids[0], loss=loss)
