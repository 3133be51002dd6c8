"""Vendors the UNMODIFIED hot-path files of the reference into oracle/_ref/ (git-ignored, travels to the GPU box).

    python oracle/build_ref.py            # needs /root/reference (read-only mount); run by __graft_entry__.build()

The reference is a pure-Python tree, so "building" it is a byte-for-byte copy of the files its quantisation path
imports (modules.quantize / rqvae / loss / normalize / encoder, init.kmeans, distributions.gumbel, data.schemas) from
where they lie under /root/reference; nothing is edited and nothing lands in git history.  `bench.py --impl reference`
and bench.py's `cpu_baseline` leg import it from there (behind the name-only `gin` stub in oracle/_stubs) and time
the reference's own `Quantize.forward` + autograd on the box's host cores (cpu_baseline.kind = "reference").
A manifest with the sha256 of every copied file is written next to them.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("RQ_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = [
    "modules/__init__.py", "modules/quantize.py", "modules/rqvae.py", "modules/loss.py", "modules/normalize.py",
    "modules/encoder.py", "init/kmeans.py", "distributions/gumbel.py", "data/__init__.py", "data/schemas.py",
]


def available() -> bool:
    return os.path.exists(os.path.join(DST, "MANIFEST.json"))


def build(force: bool = False) -> bool:
    """Returns True when oracle/_ref is in place (copied now or earlier)."""
    if not os.path.isdir(os.path.join(SRC, "modules")):
        return available()
    manifest = {}
    for rel in FILES:
        src = os.path.join(SRC, rel)
        if not os.path.exists(src):
            continue
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    json.dump({"source": SRC, "files": manifest}, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1)
    return True


def load():
    """Imports the vendored reference (oracle/_ref) — bench.py's reference arm only."""
    import types
    stubs = os.path.join(HERE, "_stubs")
    for p in (DST, stubs):
        if p in sys.path:
            sys.path.remove(p)
    sys.path.insert(0, DST)
    sys.path.insert(0, stubs)
    for name in [m for m in sys.modules if m.split(".")[0] in ("modules", "init", "distributions", "data")]:
        del sys.modules[name]
    import modules.quantize as q
    import init.kmeans as km
    return types.SimpleNamespace(quantize=q, kmeans=km)


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    print("oracle/_ref", "ready" if ok else "unavailable (no reference mount)")
