"""Real multi-GPU parity, one process per GPU under torchrun over NCCL: the KV-sharded decode (both the
all-gather form and the exchange fused into the merge kernel over NVLink peer memory) against the fp32 oracle on the
full, unsharded problem, and the (batch, kv-head)-sharded prefill against the oracle.  Needs >= 2 CUDA devices
(`gpurun --gpus 2 -- python -m pytest tests -m multigpu`); skipped on a single-GPU box.  The CPU-side arithmetic of the
same path runs under gloo in tests/test_parallel_cpu.py."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_paths_under_torchrun(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} CUDA devices, have {torch.cuda.device_count()}")
    port = 29600 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "multi_gpu_check.py"), "--parity-only"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    checks = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{") and '"check"' in l]
    names = {c["check"] for c in checks}
    assert {"decode_kv_sharded parity", "decode_kv_sharded FUSED exchange parity", "prefill_sharded parity"} <= names
    for c in checks:
        assert c["ok"] and c["world"] == world and c["max_abs_err"] <= c["tol"], c
