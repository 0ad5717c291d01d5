"""Pin the oracle (CPU, no GPU needed):

* the C restatement (oracle/attention_oracle.c) against the fp32 PyTorch oracle the reference's own tests use;
* both against golden outputs of the UNMODIFIED reference CUDA kernels (tests/golden/*.npz, produced on a B200
  by tests/golden/make_golden.py from oracle/_ref).
"""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import ref_attention as R

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _t(a, bf16=False):
    t = torch.from_numpy(np.asarray(a))
    if t.dtype in (torch.float32, torch.float16):
        return t.to(torch.bfloat16 if bf16 else torch.float16)
    return t


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("shape", [(2, 100, 100, 4, 2, 64), (1, 70, 130, 2, 2, 128), (1, 1, 1, 1, 1, 32)])
def test_c_oracle_matches_torch_oracle_prefill(shape, causal):
    B, Sq, Sk, H, Hkv, D = shape
    torch.manual_seed(0)
    q, k, v = torch.randn(B, Sq, H, D).half(), torch.randn(B, Sk, Hkv, D).half(), torch.randn(B, Sk, Hkv, D).half()
    a = R.attention_ref(q, k, v, causal)
    b = R.c_prefill(q, k, v, causal)
    assert (a - b).abs().max().item() < 2e-3  # only difference: P rounded to fp16 before PV (prefill.cuh:555-574)


def test_c_oracle_matches_torch_oracle_varlen_paged():
    torch.manual_seed(1)
    page, nb = 16, 40
    kp, vp = torch.randn(nb, page, 2, 64).half(), torch.randn(nb, page, 2, 64).half()
    bt = torch.randperm(nb)[:36].reshape(3, 12).int()
    cq = torch.tensor([0, 5, 6, 30], dtype=torch.int32)
    ck = torch.tensor([0, 100, 290, 470], dtype=torch.int32)
    q = torch.randn(30, 8, 64).half()
    a = R.varlen_ref(q, kp, vp, cq, ck, True, bt)
    b = R.c_prefill(q, kp, vp, True, cq, ck, bt)
    assert (a - b).abs().max().item() < 2e-3


@pytest.mark.parametrize("splits", [1, 2, 5])
def test_c_oracle_matches_torch_oracle_decode(splits):
    torch.manual_seed(2)
    q = torch.randn(3, 1, 8, 128).half()
    kc, vc = torch.randn(3, 500, 2, 128).half(), torch.randn(3, 500, 2, 128).half()
    sl = torch.tensor([500, 257, 1], dtype=torch.int32)
    a, la = R.kvcache_ref(q, kc, vc, sl, return_lse=True)
    b, lb = R.c_decode(q, kc, vc, sl, num_splits=splits, return_lse=True)
    assert (a - b).abs().max().item() < 1e-5
    assert (la - lb).abs().max().item() < 1e-5


def test_merge_partials_is_associative_with_full_attention():
    torch.manual_seed(3)
    q = torch.randn(2, 1, 4, 64).half()
    kc, vc = torch.randn(2, 300, 4, 64).half(), torch.randn(2, 300, 4, 64).half()
    full, lse = R.kvcache_ref(q, kc, vc, None, return_lse=True)
    parts = [R.kvcache_ref(q, kc[:, a:b], vc[:, a:b], None, return_lse=True) for a, b in ((0, 100), (100, 101), (101, 300))]
    o, l = R.merge_partials_ref([p[0][:, 0] for p in parts], [p[1] for p in parts])
    assert (o - full[:, 0]).abs().max().item() < 1e-5
    assert (l - lse).abs().max().item() < 1e-5


def test_golden_fixtures_exist():
    assert len(GOLDEN) >= 8, "tests/golden/*.npz missing: run tests/golden/make_golden.py on a B200 box"


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracles_match_reference_golden(path):
    """Reference CUDA output (fp16/bf16) vs both oracles on the stored inputs.  Tolerance: the reference's own
    test tolerance (max-abs < 0.01, tests/test_mha.py:55-91) -- in practice we see <= 2e-3 (fp16) / 1e-2 (bf16)."""
    z = np.load(path)
    kind, bf16 = str(z["kind"]), bool(z["bf16"])
    tol = 1e-2 if bf16 else 2e-3
    q, k, v = _t(z["q"], bf16), _t(z["k"], bf16), _t(z["v"], bf16)
    out = torch.from_numpy(z["out"]).float()
    if kind == "prefill":
        causal = bool(z["causal"])
        a, b = R.attention_ref(q, k, v, causal), R.c_prefill(q, k, v, causal)
    elif kind == "varlen":
        cq, ck = torch.from_numpy(z["cu_q"]), torch.from_numpy(z["cu_k"])
        a, b = R.varlen_ref(q, k, v, cq, ck, bool(z["causal"])), R.c_prefill(q, k, v, bool(z["causal"]), cq, ck)
    elif kind == "varlen_paged":
        cq, ck, bt = torch.from_numpy(z["cu_q"]), torch.from_numpy(z["cu_k"]), torch.from_numpy(z["block_table"])
        a = R.varlen_ref(q, k, v, cq, ck, bool(z["causal"]), bt)
        b = R.c_prefill(q, k, v, bool(z["causal"]), cq, ck, bt)
    else:
        bt, sl = torch.from_numpy(z["block_table"]), torch.from_numpy(z["seqlens"])
        a = R.kvcache_ref(q, k, v, sl, bt)
        b = R.c_decode(q, k, v, sl, bt, num_splits=int(z["splits"]))
    assert (a - out).abs().max().item() <= tol, "torch fp32 oracle vs reference CUDA kernels"
    assert (b - out).abs().max().item() <= tol, "C restatement vs reference CUDA kernels"
