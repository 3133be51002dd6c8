"""Import the UNMODIFIED reference (AdamLTy/RQ-VAE-Recommender) from /root/reference.

TEST INFRASTRUCTURE, container-only.  /root/reference does not exist on the GPU
box, so nothing under tests/ -m gpu, smoke() or bench.py may import this module.
It is used by oracle/make_golden.py (fixture generation) and by the optional
`test_oracle_vs_live_reference` test, which skips when the mount is absent.

gin-config / polars / torch_geometric / sentence_transformers are not installed
and there is no network; oracle/_stubs holds name-only stand-ins.
"""
import os
import sys

REFERENCE_ROOT = os.environ.get("RQ_REFERENCE_ROOT", "/root/reference")
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_stubs")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "modules"))


def load():
    """Returns a namespace with the reference's hot-path modules."""
    if not available():
        raise RuntimeError(f"reference not mounted at {REFERENCE_ROOT}")
    for p in (REFERENCE_ROOT, _STUBS):
        if p in sys.path:
            sys.path.remove(p)
    # stubs first so that `import gin` resolves, then the reference's own tree
    sys.path.insert(0, REFERENCE_ROOT)
    sys.path.insert(0, _STUBS)
    # the reference calls torch.set_float32_matmul_precision('high') at import
    # (modules/rqvae.py:19); on CPU this is bit-identical to 'highest'.
    import types
    import modules.quantize as q
    import modules.rqvae as r
    import modules.loss as lo
    import modules.normalize as nz
    import init.kmeans as km
    import distributions.gumbel as gb
    import modules.tokenizer.semids as sem
    import data.schemas as schemas
    return types.SimpleNamespace(quantize=q, rqvae=r, loss=lo, normalize=nz, kmeans=km,
                                 gumbel=gb, semids=sem, schemas=schemas)
