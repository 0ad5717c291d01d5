#!/usr/bin/env python
"""Tuning aid (GPU box): per-role clock64() timeline of prefill CTA (0,0,0) from a -DMFA_TRACE build.
    MFA_B200_LIB=.../libtrace.so python scripts/trace_prefill.py [S]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import statistics
import torch
import mini_flash_attention as mfa
from mini_flash_attention_b200 import _lib
S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
CAUSAL = len(sys.argv) > 2 and sys.argv[2] == "causal"
B, H, D = 48, 24, 128
q, k, v = (torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16) for _ in range(3))
buf = torch.zeros(4 * 64 * 8 + 64 * 8, dtype=torch.int64, device="cuda")
for _ in range(3):
    mfa.flash_attn_func(q, k, v, causal=CAUSAL)
_lib.load().mfa_debug_set_trace(buf.data_ptr())
mfa.flash_attn_func(q, k, v, causal=CAUSAL)
torch.cuda.synchronize()
_lib.load().mfa_debug_set_trace(None)
ct = buf.cpu()[4 * 64 * 8:].view(64, 8)
pr = buf.cpu()[3 * 64 * 8:4 * 64 * 8].view(64, 8)
t = buf.cpu()[:3 * 64 * 8].view(3, 64, 8)
t0 = int(t[t > 0].min())
r = lambda x: int(x) - t0 if int(x) > 0 else -1
print("softmax events: 0 loop top, 1 S ready, 2 S in regs, 6 row max done, 7 grow vote done, 3 rescale done, 4 P_lo released, 5 P_hi released")
print("mma events (single): 0 K/V ready, 1 p_lo0 seen, 2 p_hi0 seen, 3 PV0+S0 issued, 4 p_lo1 seen, 5 p_hi1 seen, 6 PV1+S1 issued")
n = min(S // 128, 64)
if CAUSAL:
    # causal: CTA 0 of a 147-CTA grid (coprime to the m-block count) runs items 0, 147, 294, ...; m-block = nmb-1 - item % nmb;
    # Q tile i of m-block mb has 2 mb + 1 + i KV tiles
    ep = buf.cpu()[3 * 64 * 8:5 * 64 * 8].view(2, 64, 8)
    nmb = S // 256
    G = 148
    import math
    while nmb > 1 and math.gcd(G, nmb) != 1: G -= 1
    base = [0, 0]
    mbase = 0
    print(f"causal item timeline, S={S}, grid {G}: per item  m-block | MMA Q0+K landed | WGi: loop top, first S ready, last P released, o_full seen, epilogue done")
    k = 0
    while True:
        item = k * G
        mb = nmb - 1 - item % nmb
        n = [2 * mb + 1, 2 * mb + 2]
        if max(base[0] + n[0], base[1] + n[1], mbase + n[1]) > 64 or k >= 24: break
        line = f"  item {k:2d} mb {mb:2d} prod {[r(ep[0, k, e]) for e in (4, 5, 6, 7)]} MMA start {r(t[2, mbase, 7])}"
        for i in (0, 1):
            line += f"  WG{i} [{r(t[i, base[i], 0])}, {r(t[i, base[i], 1])}, {r(t[i, base[i] + n[i] - 1, 5])}, {r(ep[i, k, 1])}, {r(ep[i, k, 2])}]"
        print(line)
        if mb == nmb - 1 and k > 0:
            print("    steps of this item (relative to its MMA start): MMA [K/V ready, p_lo0, p_hi0, PV0+S0 issued, p_lo1, p_hi1, PV1+S1 issued] | WG0 [S ready, max done, P_lo, P_hi] | WG1 [...]")
            z = int(t[2, mbase, 7])
            rr = lambda x: int(x) - z if int(x) > 0 else -1
            for j in range(n[1]):
                l2 = f"      j={j} MMA {[rr(t[2, mbase + j, e]) for e in range(7)]}"
                if j < n[0]: l2 += f" | WG0 {[rr(t[0, base[0] + j, e]) for e in (1, 6, 4, 5)]}"
                l2 += f" | WG1 {[rr(t[1, base[1] + j, e]) for e in (1, 6, 4, 5)]}"
                print(l2)
        base[0] += n[0]; base[1] += n[1]; mbase += n[1]; k += 1
    sys.exit(0)
if S <= 2048:
    # short sequences: the trace runs over the first 64 KV steps of CTA 0, i.e. several items; show the item boundaries
    ep = buf.cpu()[3 * 64 * 8:5 * 64 * 8].view(2, 64, 8)
    print(f"item timeline (S={S}: {n} KV steps per item); all times relative to the first stamp")
    print("  prod: reaches item / Q issued / first K issued / last load issued;  MMA: Q+K landed;  WGi: item loop top, first S ready, last P released, o_full seen, epilogue done")
    for k in range(min(64 // n, 12)):
        b = k * n
        line = f"  item {k:2d} prod {[r(ep[0, k, e]) for e in (4, 5, 6, 7)]} MMA start {r(t[2, b, 7])}"
        for i in (0, 1):
            line += f"  WG{i} [{r(t[i, b, 0])}, {r(t[i, b, 1])}, {r(t[i, b + n - 1, 5])}, {r(ep[i, k, 1])}, {r(ep[i, k, 2])}]"
        print(line)
    for i in (0, 1):
        d = [int(t[i, (k + 1) * n, 1] - t[i, k * n, 1]) for k in range(1, min(64 // n, 12) - 1)]
        print(f"  WG{i} item period median {statistics.median(d)}  (= {n} steps x {statistics.median(d) / n:.0f}); steady-state step from long sequences ~3440")
        gap = [int(t[i, (k + 1) * n, 1] - t[i, k * n + n - 1, 5]) for k in range(1, min(64 // n, 12) - 1)]
        print(f"  WG{i} last P released -> next item's first S ready: median {statistics.median(gap)}")
    sys.exit(0)
for j in list(range(0, 6)) + list(range(n - 3, n)):
    print(f"j={j:2d} WG0 {[r(x) for x in t[0, j, :8]]}  WG1 {[r(x) for x in t[1, j, :8]]}  MMA {[r(x) for x in t[2, j, :8]]}")
import statistics
for role, ev in ((0, 1), (1, 1), (2, 0)):
    d = [int(t[role, j + 1, ev] - t[role, j, ev]) for j in range(4, n - 2)]
    print("role", role, "period median", statistics.median(d), "min", min(d), "max", max(d))
for role in (0, 1):
    busy = [int(t[role, j, 5] - t[role, j, 1]) for j in range(4, n - 1)]
    wait = [int(t[role, j + 1, 1] - t[role, j, 5]) for j in range(4, n - 2)]
    order = [1, 2, 6, 7, 3, 4, 5]
    seg = [[int(t[role, j, order[e + 1]] - t[role, j, order[e]]) for j in range(4, n - 1)] for e in range(len(order) - 1)]
    print("WG", role, "busy median", statistics.median(busy), "wait-for-S median", statistics.median(wait),
          "segments ld / max / vote / rescale / half0 / half1", [statistics.median(s) for s in seg])
