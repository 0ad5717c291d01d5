import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import mini_flash_attention as mfa
from oracle import ref_attention as R
from test_varlen_gpu import _paged_inputs, _cu
for (sq, sk, H, Hkv, D, page, causal, poison) in [
    ([50, 260], [333, 260], 4, 2, 128, 8, False, True),
    ([50, 260], [333, 260], 4, 2, 128, 8, False, False),
    ([50, 260], [333, 260], 4, 2, 128, 16, False, True),
    ([50, 260], [333, 260], 4, 2, 128, 32, False, True),
    ([50], [333], 1, 1, 128, 8, False, False),
    ([50], [64], 1, 1, 128, 8, False, False),
    ([50], [128], 1, 1, 128, 8, False, False),
    ([50], [256], 1, 1, 128, 8, False, False),
]:
    q, kc, vc, bt = _paged_inputs(sq, sk, H, Hkv, D, page, torch.float16, 0, poison=poison)
    cu_q, cu_k = _cu(sq), _cu(sk)
    out = mfa.flash_attn_varlen_func(q, kc, vc, cu_q, cu_k, max(sq), max(sk), causal=causal, block_table=bt)
    ref = R.varlen_ref(q, torch.nan_to_num(kc), torch.nan_to_num(vc), cu_q, cu_k, causal, bt)
    err = (out.float() - ref).abs()
    print(sq, sk, H, Hkv, page, poison, "max err", err.max().item(), "per-seq", [err[cu_q[i]:cu_q[i+1]].max().item() for i in range(len(sq))],
          "rows>tol", (err.amax(dim=(1, 2)) > 2e-3).nonzero().flatten().tolist()[:10], "heads", err.amax(dim=(0, 2)).tolist())
