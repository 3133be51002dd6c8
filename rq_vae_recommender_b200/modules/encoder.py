"""Bias-free SiLU MLP (reference: modules/encoder.py:7-36); state-dict keys (`mlp.{0,2,4,6}.weight`) are the reference's.

Training (autograd) runs the nn.Sequential on torch/cuBLAS like the reference.  Inference on a CUDA device — the
tokenisation path `RqVae.get_semantic_ids` -> `encode` (modules/rqvae.py:117-118, 181-183) — runs every Linear through
the engine's encoder-layer kernel (`rq_encoder_layer`: TMA-fed TF32 tcgen05 GEMM with SiLU / the L2 normalisation fused
into its epilogue) when the process allows TF32 matmuls, i.e. `torch.get_float32_matmul_precision() != "highest"` — the
setting the reference itself makes (`set_float32_matmul_precision('high')`, modules/rqvae.py:19) and the same switch
that lets torch pick TF32 for its own cuBLAS calls.  `RQ_ENCODER_TORCH=1` keeps torch everywhere."""
import os
from typing import List

import torch
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

    # ---- engine path ---------------------------------------------------------------------------------------
    def _engine_plan(self, x: Tensor):
        """[(weight, act flags)] when this call can run on the engine, else None."""
        if (not x.is_cuda or x.dtype != torch.float32 or x.dim() != 2 or x.shape[0] == 0
                or torch.get_float32_matmul_precision() == "highest" or os.environ.get("RQ_ENCODER_TORCH") == "1"
                or torch.compiler.is_compiling()):
            return None
        if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            return None
        from .. import functional as RF
        plan, mods = [], list(self.mlp)
        i = 0
        while i < len(mods):
            m = mods[i]
            if isinstance(m, nn.Identity):
                i += 1
                continue
            if not isinstance(m, nn.Linear) or m.bias is not None or m.weight.dtype != torch.float32:
                return None
            act, j = RF.ACT_NONE, i + 1
            if j < len(mods) and isinstance(mods[j], nn.SiLU):
                act, j = RF.ACT_SILU, j + 1
                if j < len(mods) and isinstance(mods[j], nn.Dropout):
                    if self.training and mods[j].p > 0:
                        return None
                    j += 1
            elif j < len(mods) and isinstance(mods[j], L2NormalizationLayer):
                if mods[j].dim not in (-1, 1) or mods[j].eps != 1e-12 or m.out_features > 64:
                    return None
                act, j = RF.ACT_L2NORM, j + 1
            plan.append((m.weight, act))
            i = j
        for k, (w, act) in enumerate(plan):
            flags = act | (RF.ACT_TF32_OUT if k + 1 < len(plan) else 0)     # hidden activations feed the next product
            if not RF.encoder_layer_supported(x.shape[0], w.shape[1], w.shape[0], flags):
                return None
            plan[k] = (w, flags)
        return plan

    def _tf32_weight(self, k: int, w: Tensor) -> Tensor:
        """The layer's weight rounded to nearest onto the TF32 grid (the tensor core would truncate it), cached until the
        parameter changes."""
        cache = self.__dict__.setdefault("_tf32_weights", {})
        key = (w.data_ptr(), w._version, w.device)
        hit = cache.get(k)
        if hit is None or hit[0] != key:
            bits = w.detach().contiguous().view(torch.int32)
            hit = (key, ((bits + 0x1000) & ~0x1FFF).view(torch.float32))
            cache[k] = hit
        return hit[1]

    def forward(self, x: Tensor) -> Tensor:
        assert x.shape[-1] == self.input_dim, f"Invalid input dim: Expected {self.input_dim}, found {x.shape[-1]}"
        plan = self._engine_plan(x)
        if plan is None:
            return self.mlp(x)
        from .. import functional as RF
        h = x
        for k, (w, flags) in enumerate(plan):
            h = RF.encoder_layer(h, self._tf32_weight(k, w), flags)
        return h
