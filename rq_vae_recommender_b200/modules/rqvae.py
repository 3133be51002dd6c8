"""RQ-VAE behind the reference's `RqVae` interface (modules/rqvae.py:22-165).

Same constructor, attributes, NamedTuple outputs, memory layouts and state-dict keys
(`layers.{i}.embedding.weight`, `layers.{i}.out_proj.0.weight`, `encoder.mlp.*`, `decoder.mlp.*`).
The level loop of `get_semantic_ids` (:124-131) is ONE fused CUDA launch over all L levels; the
encoder / decoder MLPs stay on torch (cuBLAS), outside the named path.
"""
import os
import warnings
import weakref
from functools import cached_property
from typing import List, NamedTuple

import torch
from torch import Tensor, nn

from .. import functional as RF
from ..data.schemas import SeqBatch
from .encoder import MLP
from .loss import CategoricalReconstuctionLoss, ReconstructionLoss
from .normalize import l2norm
from .quantize import GumbelNoise, Quantize, QuantizeForwardMode, kernel_distance, kernel_mode

try:  # the reference mixes in huggingface_hub.PyTorchModelHubMixin (rqvae.py:37)
    from huggingface_hub import PyTorchModelHubMixin
except Exception:  # pragma: no cover
    class PyTorchModelHubMixin:  # type: ignore
        pass


_COMPILED: "weakref.WeakKeyDictionary" = weakref.WeakKeyDictionary()   # RqVae instance -> compiled forward + status


class _Placeholder:
    """Stands in for a class of the reference tree that is not importable while reading its checkpoint."""

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        pass

    def __call__(self, *a, **k):          # enum lookups such as QuantizeForwardMode(3)
        return self


def load_reference_checkpoint(path: str, map_location=None) -> dict:
    """torch.load of a reference checkpoint; falls back to a tolerant unpickler when the file refers to classes of
    the reference's own module tree (`modules.*`, `init.*`, ...) that cannot be imported here."""
    try:
        return torch.load(path, map_location=map_location, weights_only=False)
    except (ModuleNotFoundError, AttributeError, ImportError):
        pass
    import pickle
    import types

    class _TolerantUnpickler(pickle.Unpickler):
        def find_class(self, module, name):
            try:
                return super().find_class(module, name)
            except (ModuleNotFoundError, AttributeError, ImportError):
                return type(name, (_Placeholder,), {"__module__": module})

    shim = types.ModuleType("rq_tolerant_pickle")
    shim.Unpickler = _TolerantUnpickler
    shim.load = lambda f, **kw: _TolerantUnpickler(f, **kw).load()
    shim.__name__ = "pickle"
    for attr in ("dumps", "dump", "loads", "Pickler", "PickleError", "UnpicklingError", "HIGHEST_PROTOCOL",
                 "DEFAULT_PROTOCOL"):
        setattr(shim, attr, getattr(pickle, attr))
    return torch.load(path, map_location=map_location, weights_only=False, pickle_module=shim)


class RqVaeOutput(NamedTuple):
    embeddings: Tensor
    residuals: Tensor
    sem_ids: Tensor
    quantize_loss: Tensor


class RqVaeComputedLosses(NamedTuple):
    loss: Tensor
    reconstruction_loss: Tensor
    rqvae_loss: Tensor
    embs_norm: Tensor
    p_unique_ids: Tensor


class RqVae(nn.Module, PyTorchModelHubMixin):
    def __init__(
        self,
        input_dim: int,
        embed_dim: int,
        hidden_dims: List[int],
        codebook_size: int,
        codebook_kmeans_init: bool = True,
        codebook_normalize: bool = False,
        codebook_sim_vq: bool = False,
        codebook_mode: QuantizeForwardMode = QuantizeForwardMode.GUMBEL_SOFTMAX,
        n_layers: int = 3,
        commitment_weight: float = 0.25,
        n_cat_features: int = 18,
    ) -> None:
        self._config = locals()
        super().__init__()
        self.input_dim = input_dim
        self.embed_dim = embed_dim
        self.hidden_dims = hidden_dims
        self.n_layers = n_layers
        self.codebook_size = codebook_size
        self.commitment_weight = commitment_weight
        self.n_cat_feats = n_cat_features
        self.layers = nn.ModuleList(modules=[
            Quantize(embed_dim=embed_dim, n_embed=codebook_size, forward_mode=codebook_mode,
                     do_kmeans_init=codebook_kmeans_init,
                     codebook_normalize=(i == 0 and codebook_normalize),      # level 0 only, rqvae.py:70
                     sim_vq=codebook_sim_vq, commitment_weight=commitment_weight)
            for i in range(n_layers)
        ])
        self.encoder = MLP(input_dim=input_dim, hidden_dims=hidden_dims, out_dim=embed_dim,
                           normalize=codebook_normalize)
        self.decoder = MLP(input_dim=embed_dim, hidden_dims=hidden_dims[-1::-1], out_dim=input_dim,
                           normalize=True)
        self.reconstruction_loss = (CategoricalReconstuctionLoss(n_cat_features) if n_cat_features != 0
                                    else ReconstructionLoss())

    @cached_property
    def config(self) -> dict:
        return self._config

    @property
    def device(self) -> torch.device:
        return next(self.encoder.parameters()).device

    def load_pretrained(self, path: str) -> None:
        """Loads a checkpoint written by the reference's training driver (train_rqvae.py:209-221:
        {"iter", "model", "model_config", "optimizer"}; rqvae.py:103-106).  `model_config` is `RqVae.config`, i.e. the
        `locals()` of the reference's `__init__`: it pickles the reference's own classes (`modules.rqvae.RqVae`,
        `modules.quantize.QuantizeForwardMode`).  When those modules are not importable the file is read again with
        an unpickler that stands unknown classes in with inert placeholders; only "model" and "iter" are used."""
        state = load_reference_checkpoint(path, map_location=self.device)
        self.load_state_dict(state["model"])
        print(f"---Loaded RQVAE Iter {state['iter']}---")

    def encode(self, x: Tensor) -> Tensor:
        return self.encoder(x)

    def decode(self, x: Tensor) -> Tensor:
        return self.decoder(x)

    # ---------------------------------------------------------------------------------------
    def effective_codebooks(self) -> Tensor:
        """[L, K, D] stack of `out_proj(embedding.weight)`; autograd carries the gradient of the
        fused kernel back to every level's parameters (incl. sim_vq / normalisation)."""
        return torch.stack([layer.effective_codebook() for layer in self.layers])

    def _uniform_config(self) -> bool:
        l0 = self.layers[0]
        return all(l.forward_mode == l0.forward_mode and l.distance_mode == l0.distance_mode
                   and l.quantize_loss.commitment_weight == l0.quantize_loss.commitment_weight
                   and l.n_embed == l0.n_embed and l.training == l0.training for l in self.layers)

    def quantize_latents(self, res: Tensor, gumbel_t: float = 0.001) -> RqVaeOutput:
        """Everything of get_semantic_ids after the encoder (rqvae.py:121-138)."""
        if any(layer.needs_kmeans_init() for layer in self.layers) or not self._uniform_config():
            return self._quantize_latents_per_level(res, gumbel_t)
        l0 = self.layers[0]
        mode = kernel_mode(l0.forward_mode, l0.training)
        gu, seed, offset = (None, 0, 0)
        if mode == RF.MODE_GUMBEL_SOFTMAX:
            gu, seed, offset = GumbelNoise.next(self.n_layers, res.shape[0], self.codebook_size, res.device)
        emb, residuals, ids, loss = RF.residual_quantize(
            res, self.effective_codebooks(), mode, float(l0.quantize_loss.commitment_weight), float(gumbel_t),
            kernel_distance(l0.distance_mode), gu, seed, offset)
        # the reference returns [B,D,L] / [B,L] VIEWS over [L,B,D] / [L,B] memory (einops rearrange of
        # a stacked list, rqvae.py:133-136); same strides here
        return RqVaeOutput(embeddings=emb.permute(1, 2, 0), residuals=residuals.permute(1, 2, 0),
                           sem_ids=ids.permute(1, 0), quantize_loss=loss)

    def _quantize_latents_per_level(self, res: Tensor, gumbel_t: float) -> RqVaeOutput:
        """Level-at-a-time path: needed while codebooks are being k-means-initialised, because
        level l's k-means input is the residual left by the freshly initialised levels < l
        (quantize.py:102-103, rqvae.py:124-131)."""
        quantize_loss = 0
        embs, residuals, sem_ids = [], [], []
        for layer in self.layers:
            residuals.append(res)
            quantized = layer(res, temperature=gumbel_t)
            quantize_loss = quantize_loss + quantized.loss
            res = res - quantized.embeddings
            sem_ids.append(quantized.ids)
            embs.append(quantized.embeddings)
        return RqVaeOutput(embeddings=torch.stack(embs).permute(1, 2, 0),
                           residuals=torch.stack(residuals).permute(1, 2, 0),
                           sem_ids=torch.stack(sem_ids).permute(1, 0), quantize_loss=quantize_loss)

    def get_semantic_ids(self, x: Tensor, gumbel_t: float = 0.001) -> RqVaeOutput:
        return self.quantize_latents(self.encode(x), gumbel_t)

    # The reference decorates forward with @torch.compile(mode="reduce-overhead") (rqvae.py:140).  Same here, through
    # the torch.library ops of ops.py (Dynamo cannot trace ctypes): Inductor fuses the encoder / decoder / loss
    # elementwise work, the fused quantiser is one opaque node forward and one backward, and the whole step replays
    # as CUDA graphs.  Class-level switch: `RqVae.compile_forward = False` or RQ_COMPILE_FORWARD=0 keeps it eager.
    compile_forward: bool = os.environ.get("RQ_COMPILE_FORWARD", "1") != "0"
    compile_mode: str = "reduce-overhead"

    def _wants_compiled(self, x: Tensor) -> bool:
        if not (self.compile_forward and x.is_cuda) or torch.compiler.is_compiling():
            return False
        if _COMPILED.get(self, {}).get("failed", False):
            return False
        # the first training forward runs k-means on the host side of every level (data-dependent Python): eager
        return not any(layer.needs_kmeans_init() for layer in self.layers) and self._uniform_config()

    def forward(self, batch: SeqBatch, gumbel_t: float) -> RqVaeComputedLosses:
        """Training losses (rqvae.py:141-165)."""
        if self._wants_compiled(batch.x):
            st = _COMPILED.setdefault(self, {})              # kept outside the module: state dict / deepcopy unchanged
            if "fn" not in st:
                st["fn"] = torch.compile(self._forward_impl, mode=self.compile_mode)
            try:
                out = st["fn"](batch, gumbel_t)
            except Exception as e:                           # e.g. no compiler toolchain on this host: stay eager
                if st.get("ok", False):
                    raise                                    # it worked before: a real error, not a set-up problem
                st["failed"] = True
                warnings.warn(f"RqVae.forward: torch.compile failed ({type(e).__name__}: {e}); running eager")
                return self._forward_impl(batch, gumbel_t)
            st["ok"] = True
            return out
        return self._forward_impl(batch, gumbel_t)

    def _forward_impl(self, batch: SeqBatch, gumbel_t: float) -> RqVaeComputedLosses:
        x = batch.x
        quantized = self.get_semantic_ids(x, gumbel_t)
        embs = quantized.embeddings
        x_hat = self.decode(embs.sum(axis=-1))
        if self.n_cat_feats != 0:
            x_hat = torch.cat([l2norm(x_hat[..., :-self.n_cat_feats]), x_hat[..., -self.n_cat_feats:]], axis=-1)
        else:
            # x_hat[..., :-0] is EMPTY in the reference (rqvae.py:146): with n_cat_features == 0 the
            # decoder output passes through un-normalised (its MLP already ends in an L2 norm)
            x_hat = torch.cat([l2norm(x_hat[..., :0]), x_hat], axis=-1)
        reconstruction_loss = self.reconstruction_loss(x_hat, x)
        rqvae_loss = quantized.quantize_loss
        loss = (reconstruction_loss + rqvae_loss).mean()
        with torch.no_grad():
            embs_norm = embs.norm(dim=1)
            # fraction of rows with no later duplicate == number of distinct tuples / B (rqvae.py:155-157);
            # computed with the sort-based dedup kernel instead of a [B,B,L] comparison
            table = RF.corpus_table(quantized.sem_ids, self.codebook_size)
            p_unique_ids = (table[:, -1] == 0).sum() / quantized.sem_ids.shape[0]
        return RqVaeComputedLosses(loss=loss, reconstruction_loss=reconstruction_loss.mean(),
                                   rqvae_loss=rqvae_loss.mean(), embs_norm=embs_norm,
                                   p_unique_ids=p_unique_ids)
