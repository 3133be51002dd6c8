"""Optional gin-config binding.  gin is not a hard dependency: when it is absent the decorators
are no-ops, so the same modules import on a bare box ("Config / flags")."""
try:  # pragma: no cover - depends on the environment
    import gin  # type: ignore
    HAVE_GIN = True
except Exception:  # ImportError or a broken install
    gin = None
    HAVE_GIN = False


def constants_from_enum(module: str):
    """`@gin.constants_from_enum` registered under the REFERENCE's module path, so that macros such
    as `%modules.quantize.QuantizeForwardMode.ROTATION_TRICK` (configs/rqvae_amazon.gin:25) resolve
    to this package's enum."""
    def deco(cls):
        if HAVE_GIN:
            try:
                return gin.constants_from_enum(cls, module=module)
            except Exception:
                return cls
        return cls
    return deco
