#!/usr/bin/env python
"""Tuning aid (GPU box): per-role clock64() timeline of prefill CTA (0,0,0) from a -DMFA_TRACE build.
    MFA_B200_LIB=.../libtrace.so python scripts/trace_prefill.py [S]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
from mini_flash_attention_b200 import _lib
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
B, H, D = 48, 24, 128
q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16) for _ in range(3))
buf = torch.zeros(3 * 64 * 8, dtype=torch.int64, device="cuda")
for _ in range(3):
    mfa.flash_attn_func(q, k, v)
_lib.load().mfa_debug_set_trace(buf.data_ptr())
mfa.flash_attn_func(q, k, v)
torch.cuda.synchronize()
_lib.load().mfa_debug_set_trace(None)
t = buf.cpu().view(3, 64, 8)
t0 = int(t[t > 0].min())
r = lambda x: int(x) - t0 if int(x) > 0 else -1
print("softmax events: 0 loop top, 1 S ready, 2 S in regs, 3 max/rescale done, 4 P_lo released, 5 P_hi released")
print("mma events: 0 K/V ready, 1 p_lo0 seen, 2 p_hi0 seen, 3 PV0+S0 issued, 4 p_lo1 seen, 5 p_hi1 seen, 6 PV1+S1 issued")
n = min(S // 128, 64)
for j in list(range(0, 6)) + list(range(n - 3, n)):
    print(f"j={j:2d} WG0 {[r(x) for x in t[0, j, :6]]}  WG1 {[r(x) for x in t[1, j, :6]]}  MMA {[r(x) for x in t[2, j, :7]]}")
import statistics
for role, ev in ((0, 1), (1, 1), (2, 0)):
    d = [int(t[role, j + 1, ev] - t[role, j, ev]) for j in range(4, n - 2)]
    print("role", role, "period median", statistics.median(d), "min", min(d), "max", max(d))
for role in (0, 1):
    busy = [int(t[role, j, 5] - t[role, j, 1]) for j in range(4, n - 1)]
    wait = [int(t[role, j + 1, 1] - t[role, j, 5]) for j in range(4, n - 2)]
    seg = [[int(t[role, j, e + 1] - t[role, j, e]) for j in range(4, n - 1)] for e in range(1, 5)]
    print("WG", role, "busy median", statistics.median(busy), "wait-for-S median", statistics.median(wait),
          "segments ld/max/half0/half1", [statistics.median(s) for s in seg])
