"""CPU: the C-ABI library builds/loads and exports every symbol include/rq_engine.h declares.
No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "rq_engine.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rq_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def libpath():
    from rq_vae_recommender_b200 import build
    return build.build()


def test_header_symbols_exported(libpath):
    names = declared_symbols()
    assert len(names) >= 14
    lib = ctypes.CDLL(libpath)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in rq_engine.h but not exported"


def test_python_binding_covers_header(libpath):
    from rq_vae_recommender_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    l = _lib.lib()
    assert l.rq_abi_version() == 1
    assert l.rq_launch_count() == 0 or l.rq_launch_count() > 0


def test_argument_errors_without_gpu(libpath):
    """Argument validation happens before any CUDA call, so it is testable on CPU."""
    from rq_vae_recommender_b200 import _lib
    l = _lib.lib()
    # unsupported D
    rc = l.rq_quantize_forward(None, None, 8, 33, 256, 3, 0, 1, 0.25, 0.2, None, 0, 0, None, None, None, None, 0, None)
    assert rc == -2 and b"embed_dim" in l.rq_last_error()
    # bad mode / distance: the reference's messages (quantize.py:119,144)
    rc = l.rq_quantize_forward(None, None, 8, 32, 256, 3, 7, 1, 0.25, 0.2, None, 0, 0, None, None, None, None, 0, None)
    assert rc == -1 and b"Unsupported Quantize forward mode." in l.rq_last_error()
    rc = l.rq_quantize_forward(None, None, 8, 32, 256, 3, 0, 9, 0.25, 0.2, None, 0, 0, None, None, None, None, 0, None)
    assert rc == -1 and b"Unsupported Quantize distance mode." in l.rq_last_error()
    # null pointers with B > 0
    rc = l.rq_quantize_forward(None, None, 8, 32, 256, 3, 0, 1, 0.25, 0.2, None, 0, 0, None, None, None, None, 0, None)
    assert rc == -3
    # empty batch is a no-op success
    rc = l.rq_quantize_forward(None, None, 0, 32, 256, 3, 0, 1, 0.25, 0.2, None, 0, 0, None, None, None, None, 0, None)
    assert rc == 0
    assert l.rq_corpus_table_workspace_bytes(0, 3) > 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from rq_vae_recommender_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.RqEngineError):
        _lib.lib()
