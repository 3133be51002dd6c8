"""rq-engine: B200-native residual quantisation behind the module API of
AdamLTy/RQ-VAE-Recommender's quantization path (see DESIGN.md / INTEGRATION.md).

Importing this package does not load CUDA; the C-ABI library is opened on first use and a
missing library or a CPU tensor raises — there is no fallback path.
"""
from . import functional, parallel  # noqa: F401
from ._lib import LIB_PATH, RqEngineError, launch_count, lib  # noqa: F401
from .init.kmeans import Kmeans, KmeansOutput, kmeans_init_  # noqa: F401
from .modules.quantize import (GumbelNoise, Quantize, QuantizeDistance, QuantizeForwardMode,  # noqa: F401
                               QuantizeOutput)
from .modules.rqvae import RqVae, RqVaeComputedLosses, RqVaeOutput  # noqa: F401
from .modules.tokenizer.semids import SemanticIdTokenizer  # noqa: F401
from .data.schemas import SeqBatch, TokenizedSeqBatch, batch_to  # noqa: F401
from .data import utils as _data_utils  # noqa: F401  (rq.data.utils.DevicePrefetcher)
from . import data  # noqa: F401

__version__ = "0.1.0"
