#!/usr/bin/env python
"""Decode sweep (B=24, MHA H=24 and GQA 32/8, KV 512..8192, paged 256) for the library named by MFA_B200_LIB."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
dt = torch.bfloat16
for Hq, Hkv in ((24, 24), (32, 8)):
    for KV in (512, 1024, 2048, 4096, 8192):
        B, D, page = 24, 128, 256
        nb = B * KV // page
        q = torch.randn(B, 1, Hq, D, device="cuda", dtype=dt)
        kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        bt = torch.randperm(nb, device="cuda").reshape(B, KV // page).int()
        sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
        fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0)
        for _ in range(5): fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g): fn()
        for _ in range(3): g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50): g.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 50 * 1e3
        print(f"{os.environ.get('MFA_B200_LIB','default')[-16:]:>16} Hq={Hq} Hkv={Hkv} KV={KV:5d} {us:7.1f} us {2.0*B*KV*Hkv*D*2/us/1e3:7.0f} GB/s", flush=True)
