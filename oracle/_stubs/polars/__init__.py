"""Empty stand-in: the reference's data/ modules import polars at module scope."""
