#!/usr/bin/env python
"""BASELINE config 3 (varlen, GQA 32/8, causal, packed and paged): persistent snake-order walk vs one CTA per item
(MFA_PREFILL_VARLEN_NONPERSISTENT=1 in the environment)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
print("mode:", "one CTA per item" if os.environ.get("MFA_PREFILL_VARLEN_NONPERSISTENT") == "1" else "persistent, snake order")
g = torch.Generator(device="cuda").manual_seed(0)
for seqlens in ([128, 256, 512, 1024, 2048, 4096], [4096] * 6, [512] * 48, [100, 3000, 700, 4096, 33, 1500, 2048, 900]):
    cu = torch.tensor([0] + list(torch.tensor(seqlens).cumsum(0)), dtype=torch.int32, device="cuda")
    tot = sum(seqlens)
    q = torch.randn(tot, 32, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    k = torch.randn(tot, 8, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    v = torch.randn(tot, 8, 128, device="cuda", dtype=torch.bfloat16, generator=g)
    fn = lambda: mfa.flash_attn_varlen_func(q, k, v, cu, cu, max(seqlens), max(seqlens), causal=True)
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    fl = sum(4 * 32 * 128 * s * (s + 1) / 2 for s in seqlens)
    print(f"seqlens {str(seqlens)[:60]:60s} {ms:.4f} ms = {fl / ms / 1e9:.0f} TFLOP/s", flush=True)
