import os
import sys

import pytest

# tests/test_kv_exchange_gpu.py runs up to 8 "ranks" of the KV-sharded decode as 8 streams of ONE process; their kernels
# wait for each other, so no two of those streams may share a hardware work queue (the default is 8 queues per context,
# and streams created earlier in the session already occupy some).  Must be set before the CUDA context exists.  One
# process per GPU -- the real deployment -- does not need it.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

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
