#!/usr/bin/env python
"""One small workload per kernel for ncu captures (GPU box):  python scripts/profile_cases.py <case> [reps]
Cases: prefill (headline B48 H24 S4096 bf16), prefill_causal (same, causal), prefill_s512, varlen (config 3, packed),
varlen_paged, decode (B24 H24 KV8192 paged 256), decode_kv512, decode_gqa_kv1024, decode_fp8 (e4m3 cache, KV8192),
combine (B1 Hq32 Hkv8 KV=256K: split-KV + combine), append (paged KV-cache append of 1 row, B=256), xchg (1-rank fused
exchange step, KV=256K).  Every case launches its kernels `reps` times after two warm-up calls."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa

case = sys.argv[1]
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
g = torch.Generator(device="cuda").manual_seed(0)
bf = torch.bfloat16
r = lambda *s: torch.randn(*s, device="cuda", dtype=bf, generator=g)  # noqa: E731


def paged(B, KV, Hkv, D=128, page=256, dt=bf):
    nb = B * KV // page
    kc, vc = r(nb, page, Hkv, D), r(nb, page, Hkv, D)
    bt = torch.randperm(nb, device="cuda", generator=g).reshape(B, KV // page).int()
    sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
    return kc, vc, bt, sl


if case in ("prefill", "prefill_causal", "prefill_s512"):
    S = 512 if case == "prefill_s512" else 4096
    q, k, v = r(48, S, 24, 128), r(48, S, 24, 128), r(48, S, 24, 128)
    fn = lambda: mfa.flash_attn_func(q, k, v, causal=(case == "prefill_causal"))  # noqa: E731
elif case in ("varlen", "varlen_paged"):
    seqlens = [128, 256, 512, 1024, 2048, 4096]
    cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device="cuda")
    tot = sum(seqlens)
    q, k, v = r(tot, 32, 128), r(tot, 8, 128), r(tot, 8, 128)
    if case == "varlen":
        fn = lambda: mfa.flash_attn_varlen_func(q, k, v, cu, cu, 4096, 4096, causal=True)  # noqa: E731
    else:
        page = 256
        nblk = sum(-(-s // page) for s in seqlens)
        kc, vc = r(nblk, page, 8, 128), r(nblk, page, 8, 128)
        bt = torch.zeros(len(seqlens), 16, dtype=torch.int32, device="cuda")
        used = 0
        for i, s in enumerate(seqlens):
            n = -(-s // page)
            bt[i, :n] = torch.arange(used, used + n, device="cuda", dtype=torch.int32)
            used += n
        fn = lambda: mfa.flash_attn_varlen_func(q, kc, vc, cu, cu, 4096, 4096, causal=True, block_table=bt)  # noqa: E731
elif case in ("decode", "decode_kv512", "decode_gqa_kv1024", "decode_fp8"):
    Hq, Hkv, KV = (32, 8, 1024) if case == "decode_gqa_kv1024" else (24, 24, 512 if case == "decode_kv512" else 8192)
    q = r(24, 1, Hq, 128)
    kc, vc, bt, sl = paged(24, KV, Hkv)
    kw = {}
    if case == "decode_fp8":
        kc, vc = (kc * 0.25).to(torch.float8_e4m3fn), vc.to(torch.float8_e4m3fn)
        kw = dict(k_scale=4.0, v_scale=1.0)
    fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, **kw)  # noqa: E731
elif case == "combine":
    q = r(1, 1, 32, 128)
    kc, vc, bt, sl = paged(1, 262144, 8)
    fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0)  # noqa: E731
elif case == "xchg":
    from mini_flash_attention_b200 import parallel as par
    q = r(1, 1, 32, 128)
    kc, vc, bt, sl = paged(1, 262144, 8)
    x = par.KVShardExchange(1, 0, 32, 128)
    fn = lambda: par.decode_kv_sharded(q, kc, vc, sl, bt, world=1, exchange=x)  # noqa: E731
elif case == "append":
    B, Hkv, KV = 256, 8, 1024
    q = r(B, 1, 32, 128)
    kc, vc, bt, sl = paged(B, KV, Hkv)
    sl = sl - 1
    kn, vn = r(B, 1, Hkv, 128), r(B, 1, Hkv, 128)
    fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0, k=kn, v=vn)  # noqa: E731
else:
    raise SystemExit(f"unknown case {case}")
for _ in range(2):
    fn()
torch.cuda.synchronize()
for _ in range(reps):
    fn()
torch.cuda.synchronize()
print("ok", case)
