"""ctypes binding of librqengine.so (include/rq_engine.h).

PyTorch is plumbing here: it owns device memory and streams; every argument crosses the
boundary as a raw device pointer + sizes.  There is no CPU implementation behind this
module: a missing library raises, a CPU tensor raises.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_uint64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librqengine.so")

MODE_EVAL, MODE_GUMBEL_SOFTMAX, MODE_STE, MODE_ROTATION_TRICK = 0, 1, 2, 3
DIST_L2, DIST_COSINE = 1, 2
IMPL_AUTO, IMPL_SIMT, IMPL_TCGEN05 = 0, 1, 2

# every symbol include/rq_engine.h declares: (restype, argtypes)
SIGNATURES = {
    "rq_abi_version": (c_int, []),
    "rq_last_error": (c_char_p, []),
    "rq_launch_count": (c_uint64, []),
    "rq_tcgen05_supported": (c_int, [c_int] * 5),
    "rq_quantize_forward": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_float,
                                    c_float, c_void_p, c_uint64, c_uint64, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_int, c_void_p]),
    "rq_quantize_forward_ws": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_float,
                                       c_float, c_void_p, c_uint64, c_uint64, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_int, c_void_p]),
    "rq_gumbel_weights_saved": (c_int, [c_int64, c_int, c_int, c_int, c_int, c_int, c_int]),
    "rq_quantize_backward_ws": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                        c_int, c_int, c_int, c_float, c_float, c_void_p, c_uint64, c_uint64,
                                        c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "rq_quantize_backward": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                     c_int, c_int, c_int, c_float, c_float, c_void_p, c_uint64, c_uint64,
                                     c_void_p, c_void_p, c_void_p]),
    "rq_quantize_backward_ex": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                        c_int, c_int, c_int, c_float, c_float, c_void_p, c_uint64, c_uint64,
                                        c_void_p, c_void_p, c_int, c_void_p]),
    "rq_kmeans_assign_accumulate": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p, c_int, c_void_p,
                                            c_void_p, c_void_p, c_int, c_void_p]),
    "rq_kmeans_update": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                 c_void_p]),
    "rq_corpus_table_workspace_bytes": (c_int64, [c_int64, c_int]),
    "rq_corpus_table": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int64, c_void_p, c_void_p, c_int64,
                                c_void_p]),
    "rq_prefix_index_workspace_bytes": (c_int64, [c_int64]),
    "rq_prefix_index_build": (c_int, [c_void_p, c_int64, c_int, c_int, c_int64, c_void_p, c_void_p, c_int64,
                                      c_void_p]),
    "rq_prefix_exists": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p]),
    "rq_gather_tokens": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "rq_peer_allreduce_max_floats": (c_int64, []),
    "rq_peer_allreduce_state_words": (c_int, []),
    "rq_peer_allreduce_pad_bytes": (c_int, []),
    "rq_peer_allreduce_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int64, c_int64, c_void_p, c_float, c_void_p,
                                      c_void_p]),
    "rq_debug_peer_timeout_flag": (c_int, []),
    "rq_id_diversity_workspace_bytes": (c_int64, [c_int, c_int64]),
    "rq_id_diversity": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                c_void_p]),
    "rq_id_diversity_static_max_bytes": (c_int64, []),
    "rq_id_diversity_static": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "rq_debug_tc_timeout_flag": (c_int, []),
    "rq_debug_tc_set_accumulator_dump": (c_int, [c_void_p]),
    "rq_debug_tc_set_timeline": (c_int, [c_void_p]),
    "rq_debug_split_gemm": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                                    c_float, c_int, c_int, c_void_p]),
    "rq_debug_split_gemm_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int64,
                                       c_float, c_int, c_int, c_void_p]),
    "rq_encoder_layer_supported": (c_int, [c_int64, c_int, c_int, c_int]),
    "rq_encoder_layer": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "rq_semantic_ids_host": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                     c_int, c_void_p]),
}

_lib = None


class RqEngineError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Loads the library (building is `python -m rq_vae_recommender_b200.build`).  Fails loudly."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RqEngineError(
                f"{LIB_PATH} is missing: build it with `python -m rq_vae_recommender_b200.build` "
                "(there is no CPU or PyTorch fallback for the quantization path)")
        l = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        if l.rq_abi_version() != 1:
            raise RqEngineError("librqengine.so ABI version mismatch")
        _lib = l
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().rq_last_error().decode("utf-8", "replace")
        kind = {-1: "invalid argument", -2: "unsupported shape", -3: "null pointer", -4: "workspace too small",
                -5: "no device"}.get(rc, f"cuda error {rc}")
        raise RqEngineError(f"{what or 'rq-engine'}: {kind}: {msg}")


def ptr(t) -> int | None:
    return None if t is None else t.data_ptr()


def stream_of(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def require_cuda(t: torch.Tensor, name: str, dtype=None) -> torch.Tensor:
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RqEngineError(f"{name} is on {t.device}: rq-engine has no CPU path (north_star: no CPU fallback)")
    if dtype is not None and t.dtype != dtype:
        raise RqEngineError(f"{name} must be {dtype}, got {t.dtype}")
    return t


def launch_count() -> int:
    return int(lib().rq_launch_count())
