#!/usr/bin/env python
"""Split count vs step time of the paged decode at short and medium KV (input to the split heuristic, capi.cu).
Every timed launch reads a different copy of the cache (the copies together exceed the 126 MB L2), all launches of a
round are captured in one CUDA graph.    python scripts/decode_splits_gqa.py [quick]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
dt = torch.bfloat16
quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
small = len(sys.argv) > 1 and sys.argv[1] == "small"
KVS = (512, 1024, 2048, 4096, 8192)
shapes = ((24, 32, 8, KVS), (24, 24, 24, KVS)) if quick else ((24, 32, 8, KVS), (24, 24, 24, KVS), (8, 32, 8, KVS), (4, 64, 8, KVS))
if small:   # the shapes that split: few (batch x KV head) items
    shapes = ((8, 32, 8, (2048, 8192)), (4, 64, 8, (2048, 8192)), (1, 32, 8, (16384, 65536, 131072)))
for B, Hq, Hkv, kvs in shapes:
    for KV in kvs:
        D, page = 128, 256
        nb = B * KV // page
        nbytes = 2 * nb * page * Hkv * D * 2
        copies = max(2, min(16, int(600e6 // nbytes) + 1))
        q = torch.randn(B, 1, Hq, D, device="cuda", dtype=dt)
        kcs = [torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt) for _ in range(copies)]
        vcs = [torch.randn(nb, page, Hkv, D, device="cuda", dtype=dt) for _ in range(copies)]
        bt = torch.randperm(nb, device="cuda").reshape(B, KV // page).int()
        sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
        res = []
        for splits in ((0, 4, 8, 16, 37) if small else (0, 1, 2, 3, 4, 6, 8, 12)):
            if splits > KV // 64: continue
            def fn():
                for c in range(copies):
                    mfa.flash_attn_with_kvcache(q, kcs[c], vcs[c], sl, bt, splits)
            fn()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g): fn()
            for _ in range(2): g.replay()
            reps = max(3, 200 // copies)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps): g.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / (reps * copies) * 1e3
            res.append(f"{splits}:{us:.1f}")
        ideal = nbytes / 6532.2e9 * 1e6
        print(f"B={B} Hq={Hq} Hkv={Hkv} KV={KV:5d} copies={copies:2d} hbm-ideal {ideal:6.1f} us | us by splits  " + "  ".join(res), flush=True)
