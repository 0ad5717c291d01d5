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


def assert_close_16bit(ours, other, dtype_name: str, tol_scale: float = 1.0):
    """Two 16-bit implementations on the same inputs: each rounds its output once, so they may differ by one unit in
    the last place of the OUTPUT on top of the north_star bound (bf16 ulp at |o| in [2,4) is 2^-6 = 0.0156 > 1e-2: seen
    on early causal rows, which copy V almost verbatim).  |a - b| <= TOL + ulp(dtype) * |b| elementwise."""
    ulp = 2.0 ** -7 if dtype_name == "bf16" else 2.0 ** -10
    a, b = ours.float(), other.float()
    excess = ((a - b).abs() - (tol_scale * TOL[dtype_name] + ulp * b.abs())).max().item()
    assert excess <= 0.0, f"difference exceeds TOL + 1 ulp by {excess}"
