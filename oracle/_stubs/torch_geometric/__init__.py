"""Empty stand-in for torch_geometric (names only; never instantiated)."""
