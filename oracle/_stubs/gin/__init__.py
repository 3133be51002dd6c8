"""Three-symbol stand-in for gin-config (not installed, no network).

Only used by oracle/ref_import.py to import the UNMODIFIED reference from
/root/reference in the build container so that golden fixtures can be generated.
The reference touches exactly: constants_from_enum (modules/quantize.py:16,
data/processed.py:18), configurable (train_rqvae.py:24) and parse_config_file
(modules/utils.py:62).
"""


def constants_from_enum(cls=None, module=None):
    if cls is None:
        return lambda c: c
    return cls


def configurable(fn=None, **_kw):
    if fn is None or isinstance(fn, str):
        return lambda f: f
    return fn


def parse_config_file(path, *a, **k):  # pragma: no cover
    raise RuntimeError("gin stub: parse_config_file is not available")
