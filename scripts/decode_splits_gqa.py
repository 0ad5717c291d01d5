#!/usr/bin/env python
"""Split count vs step time for the B=24 GQA (32/8) decode at short KV (graph-replayed): input to the split heuristic."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
dt = torch.bfloat16
for B, Hq, Hkv in ((24, 32, 8), (8, 32, 8), (4, 64, 8)):
    for KV in (512, 1024, 2048, 4096):
        D, page = 128, 256
        nb = B * KV // page
        q = torch.randn(B, 1, Hq, D, device="cuda", dtype=dt)
        kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt)
        bt = torch.randperm(nb, device="cuda").reshape(B, KV // page).int()
        sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
        res = []
        for splits in (0, 1, 2, 3, 4, 6, 8, 12, 16):
            if splits > KV // 64: continue
            fn = lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, splits)
            for _ in range(3): fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g): fn()
            for _ in range(3): g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(50): g.replay()
            e1.record(); torch.cuda.synchronize()
            res.append(f"{splits}:{e0.elapsed_time(e1) / 50 * 1e3:.1f}")
        print(f"B={B} Hq={Hq} Hkv={Hkv} KV={KV:5d}  us by splits  " + "  ".join(res), flush=True)
