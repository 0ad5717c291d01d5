import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


@pytest.fixture(scope="session")
def mfa_lib():
    """The in-tree shared library, built on demand (nvcc cross-compiles without a GPU)."""
    from mini_flash_attention_b200 import _lib, build
    build.build()
    return _lib.load()


# tolerances stated by BASELINE.json north_star: max-abs error vs the fp32 oracle
TOL = {"fp16": 2e-3, "bf16": 1e-2}
