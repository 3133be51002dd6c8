import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# RqVae.forward compiles itself by default (the reference's @torch.compile, rqvae.py:140); the suite runs it eager
# except in tests/test_gpu_compile.py, which switches it on: one Inductor compile per model would dominate the run time
os.environ.setdefault("RQ_COMPILE_FORWARD", "0")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


def load_golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name))
