"""Batch containers at the tokenizer boundary (same fields as the reference's data/schemas.py:6-20)."""
from typing import NamedTuple, Optional

from torch import Tensor

FUT_SUFFIX = "_fut"


class SeqBatch(NamedTuple):
    user_ids: Tensor
    ids: Tensor
    ids_fut: Tensor
    x: Tensor
    x_fut: Tensor
    seq_mask: Tensor


class TokenizedSeqBatch(NamedTuple):
    user_ids: Tensor
    sem_ids: Tensor
    sem_ids_fut: Optional[Tensor]
    seq_mask: Optional[Tensor]
    token_type_ids: Tensor
    token_type_ids_fut: Tensor


def batch_to(batch: SeqBatch, device) -> SeqBatch:
    """data/utils.py:10-11 of the reference (the host->device boundary of the path)."""
    return SeqBatch(*[v.to(device) if isinstance(v, Tensor) else v for v in batch])
