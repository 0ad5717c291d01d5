#!/usr/bin/env python
"""Host-side call overhead of the public API (GPU box): per-call wall time with the GPU kept busy vs pure GPU time of
the same calls replayed from a CUDA graph.  Small decode / short prefill cases are host-bound when wall >> graph."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import mini_flash_attention as mfa
from mini_flash_attention_b200 import _C
print("host shim:", _C.SHIM)

def bench(name, fn, n=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t_issue = (time.perf_counter() - t0) / n
    torch.cuda.synchronize()
    t_wall = (time.perf_counter() - t0) / n
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(20):
                fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g.replay(); torch.cuda.synchronize()
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    print(f"{name:34s} host issue {t_issue*1e6:7.1f} us/call   wall {t_wall*1e6:7.1f} us/call   graph-replayed GPU {e0.elapsed_time(e1)*1e3/20:7.1f} us/call")

D, page = 128, 256
for (Hq, Hkv, KV) in ((24, 24, 512), (24, 24, 2048), (32, 8, 512), (32, 8, 2048)):
    B = 24
    per = KV // page; nb = B * per
    q = torch.randn(B, 1, Hq, D, device="cuda", dtype=torch.bfloat16)
    kc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
    vc = torch.randn(nb, page, Hkv, D, device="cuda", dtype=torch.bfloat16)
    bt = torch.randperm(nb, device="cuda").reshape(B, per).int()
    sl = torch.full((B,), KV, dtype=torch.int32, device="cuda")
    bench(f"decode Hq{Hq} Hkv{Hkv} KV{KV}", lambda: mfa.flash_attn_with_kvcache(q, kc, vc, sl, bt, 0))
for S in (256, 512):
    q, k, v = (torch.randn(48, S, 24, D, device="cuda", dtype=torch.bfloat16) for _ in range(3))
    bench(f"prefill S{S}", lambda: mfa.flash_attn_func(q, k, v))
